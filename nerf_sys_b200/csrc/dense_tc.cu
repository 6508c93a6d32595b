// a15: the positional-encoding MLP expert (MetaNeRF, models/inr/meta_vanilla.py:109-154) on the tensor cores.
// Unlike the 64-wide hash-grid experts its layers (63 -> 256 x8 with a skip, 42 -> 128 -> 3 by default, any
// width here) do not fit in shared memory together, so the network runs as one tcgen05 GEMM launch per layer:
//   k_dense_tc   Y = act(A W^T + b) [* relu-mask]   A, Y: fp16 "chunk-major" activations in HBM
//   k_wgrad_tc   dW = dZ^T A, db = sum dZ           accumulated in TMEM over all row tiles of a CTA
// Chunk-major = (feature chunks, ld rows, 8 halfs): the 128-row x 8-feature block a tile needs is one contiguous
// 2 KB run, read and written with one 16-byte access per thread (coalesced), and it IS the no-swizzle smem operand
// layout of tc.cuh, so tiles go global -> smem without any reshuffle.
#include "common.cuh"
#include "tc.cuh"

namespace acn {
using namespace tc;

constexpr uint32_t DROWS = 128;
constexpr uint32_t DCH = DROWS * 16;

struct DenseArgs {
  const __half* A;      // (KC chunks, ld, 8)
  int64_t ld, M;
  int KC, CT;           // input chunks; smem chunk count CT = even(1 + KC + 1): [ones | data | zero pad]
  const float* W;       // nn.Linear weight, element (n, k) at W[n*ldw + k] (transposed: W[k*ldw + n])
  const float* bias;    // (n_out) or NULL
  int64_t ldw;
  int n_out, k_valid, transposed;
  int act;              // 0 identity, 1 ReLU
  const __half* mask;   // optional: output feature n is zeroed where mask feature n <= 0 (ReLU backward)
  int64_t mask_ld;
  int mask_chunks;
  __half* out_h;        // optional chunk-major fp16 output, feature n -> chunk out_chunk0 + n/8 (< out_chunks)
  int64_t out_ld;
  int out_chunk0, out_chunks;
  float* out_f;         // optional row-major fp32 output (M, out_f_ld), columns < n_out
  int64_t out_f_ld;
};

__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

template <int NB>
__global__ void __launch_bounds__(256, 1) k_dense_tc(const __grid_constant__ DenseArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar[2];
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, q = tid >> 7;
  const uint32_t r = tid & 127;
  const int n0 = blockIdx.y * NB;
  const uint32_t wbytes = (uint32_t)a.CT * NB * 16;
  const uint32_t sw = smem_u32(smem), abuf = sw + wbytes + q * (uint32_t)a.CT * DCH;
  if (tid == 0) {
    mbar_init(smem_u32(&s_bar[0]), 1);
    mbar_init(smem_u32(&s_bar[1]), 1);
    mbar_init_fence();
  }
  __syncwarp();
  if (tid < 32) tmem_alloc(&s_tmem, NB * 2 < 32 ? 32 : NB * 2);
  {  // weight block: chunk 0 = [bias, 0..], chunks 1..KC = W (zero beyond k_valid / n_out), rest zero
    __half* d = reinterpret_cast<__half*>(smem);
    const int total = a.CT * NB * 8;
    for (int e = tid; e < total; e += 256) {
      int c = e / (NB * 8), rem = e - c * NB * 8, o = rem >> 3, j = rem & 7, n = n0 + o;
      float v = 0.f;
      if (n < a.n_out) {
        if (c == 0) {
          v = (j == 0 && a.bias) ? __ldg(a.bias + n) : 0.f;
        } else if (c <= a.KC) {
          int k = (c - 1) * 8 + j;
          if (k < a.k_valid) v = a.transposed ? __ldg(a.W + (int64_t)k * a.ldw + n) : __ldg(a.W + (int64_t)n * a.ldw + k);
        }
      }
      d[e] = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
    }
  }
  st_chunk(abuf + r * 16, 0x00003C00u, 0u, 0u, 0u);
  for (int c = a.KC + 1; c < a.CT; ++c) st_chunk(abuf + (uint32_t)c * DCH + r * 16, 0u, 0u, 0u, 0u);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = s_tmem, dcol = tbase + q * NB;
  const uint32_t trow = dcol + ((uint32_t)((tid >> 5) & 3) << 21);
  const uint32_t bar = smem_u32(&s_bar[q]);
  uint32_t phase = 0;
  constexpr uint32_t id = idesc_f16(128, NB, F16, F16, K_MAJOR, K_MAJOR);
  const int64_t ntiles = (a.M + DROWS - 1) / DROWS;
  for (int64_t tile = (int64_t)blockIdx.x * 2 + q; tile < ntiles; tile += (int64_t)gridDim.x * 2) {
    const int64_t m = tile * DROWS + r;
    const bool ok = m < a.M;
    for (int c = 0; c < a.KC; ++c) {
      uint4 v = ok ? ldg16(a.A + ((int64_t)c * a.ld + m) * 8) : make_uint4(0u, 0u, 0u, 0u);
      st_chunk(abuf + (uint32_t)(1 + c) * DCH + r * 16, v.x, v.y, v.z, v.w);
    }
    fence_before();
    fence_async_smem();
    bar_sync(1 + q, 128);
    if (r == 0) {
      fence_after();
      for (int s = 0; s < a.CT / 2; ++s)
        mma(dcol, smem_desc(abuf + s * 2 * DCH, DCH, 128), smem_desc(sw + s * 2 * (NB * 16), NB * 16, 128), id, s > 0);
      commit(bar);
    }
    mbar_wait(bar, phase);
    phase ^= 1;
    fence_after();
#pragma unroll 1
    for (int cb = 0; cb < NB / 16; ++cb) {
      float v[16];
      tmem_ld16(trow + cb * 16, v);
      tmem_ld_wait();
      const int n = n0 + cb * 16;
      if (a.act == 1) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      if (a.mask && ok) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (n / 8 + h >= a.mask_chunks) continue;
          uint4 mk = ldg16(a.mask + ((int64_t)(n / 8 + h) * a.mask_ld + m) * 8);
          const __half2* mh = reinterpret_cast<const __half2*>(&mk);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            float2 f = __half22float2(mh[i]);
            if (!(f.x > 0.f)) v[8 * h + 2 * i] = 0.f;
            if (!(f.y > 0.f)) v[8 * h + 2 * i + 1] = 0.f;
          }
        }
      }
      if (ok && a.out_h) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          int ch = n / 8 + h;
          if (ch < a.out_chunks) {
            float w[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) w[i] = fminf(fmaxf(v[8 * h + i], -65504.f), 65504.f);
            uint4 o = make_uint4(pack_h2(w[0], w[1]), pack_h2(w[2], w[3]), pack_h2(w[4], w[5]), pack_h2(w[6], w[7]));
            *reinterpret_cast<uint4*>(a.out_h + ((int64_t)(a.out_chunk0 + ch) * a.out_ld + m) * 8) = o;
          }
        }
      }
      if (ok && a.out_f) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (n + i < a.n_out) a.out_f[m * a.out_f_ld + n + i] = v[i];
      }
    }
    fence_before();
  }
  fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tbase, NB * 2 < 32 ? 32 : NB * 2);
}

// ---- weight gradient: dW (n_out, k_valid) = dZ^T A, db = column sums of dZ ------------------------------
struct WgradTcArgs {
  const __half* dZ;   // (>= ceil(n_out/64)*8 chunks readable up to dz_chunks, ld, 8)
  const __half* A;    // (KC chunks, ld, 8)
  int64_t ld, M;
  int KC, dz_chunks;
  float* partial;     // (gridDim.x, gridDim.y, 64, 8 + 8*KC)
};

__global__ void __launch_bounds__(256, 1) k_wgrad_tc(const __grid_constant__ WgradTcArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x, hf = tid >> 7;
  const uint32_t r = tid & 127;
  const uint32_t sg = smem_u32(smem), sa = sg + 8 * DCH;  // dZ tile: 8 chunks; A tile: [ones | KC chunks]
  const int NC = 8 + 8 * a.KC;                            // accumulator columns
  if (tid == 0) {
    mbar_init(smem_u32(&s_bar), 1);
    mbar_init_fence();
  }
  __syncwarp();
  if (tid < 32) tmem_alloc(&s_tmem, 512);
  if (hf == 0) st_chunk(sa + r * 16, 0x00003C00u, 0u, 0u, 0u);
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = s_tmem, bar = smem_u32(&s_bar);
  uint32_t phase = 0, acc = 0;
  const int c0 = blockIdx.y * 8;  // first dZ chunk of this CTA's 64 output features
  const int64_t ntiles = (a.M + DROWS - 1) / DROWS;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t m = tile * DROWS + r;
    const bool ok = m < a.M;
    for (int c = hf; c < 8; c += 2) {
      uint4 v = (ok && c0 + c < a.dz_chunks) ? ldg16(a.dZ + ((int64_t)(c0 + c) * a.ld + m) * 8) : make_uint4(0u, 0u, 0u, 0u);
      st_chunk(sg + (uint32_t)c * DCH + r * 16, v.x, v.y, v.z, v.w);
    }
    for (int c = hf; c < a.KC; c += 2) {
      uint4 v = ok ? ldg16(a.A + ((int64_t)c * a.ld + m) * 8) : make_uint4(0u, 0u, 0u, 0u);
      st_chunk(sa + (uint32_t)(1 + c) * DCH + r * 16, v.x, v.y, v.z, v.w);
    }
    fence_before();
    fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      fence_after();
      for (int nb = 0; nb < NC; nb += 256) {  // an MMA spans at most 256 accumulator columns
        const int nn = NC - nb < 256 ? NC - nb : 256;
        const uint32_t id = idesc_f16(64, (uint32_t)nn, F16, F16, MN_MAJOR, MN_MAJOR);
        for (int s = 0; s < 8; ++s)
          mma(tbase + nb, smem_desc(sg + s * 256, 128, DCH), smem_desc(sa + (uint32_t)(nb / 8) * DCH + s * 256, 128, DCH), id, acc | (s > 0));
      }
      commit(bar);
    }
    mbar_wait(bar, phase);  // the tiles are overwritten by the next iteration
    phase ^= 1;
    fence_after();
    acc = 1;
  }
  fence_before();
  __syncthreads();
  fence_after();
  if (tid < 128) {  // M = 64 accumulator: row o lives in lane 32*(o/16) + o%16
    const int w = tid >> 5, l = tid & 31;
    float* dst = a.partial + (((int64_t)blockIdx.x * gridDim.y + blockIdx.y) * 64 + 16 * w + l) * NC;
    float v[8];
    for (int c = 0; c < NC; c += 8) {
      tmem_ld8(tbase + c + ((uint32_t)w << 21), v);
      tmem_ld_wait();
      if (l < 16) {
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[c + i] = acc ? v[i] : 0.f;
      }
    }
  }
  fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tbase, 512);
}

__global__ void __launch_bounds__(256) k_wgrad_tc_reduce(const float* __restrict__ partial, int gx, int gy, int NC, int n_out, int k_valid,
                                                         float* __restrict__ dW, int64_t ldw, float* __restrict__ db,
                                                         const float* __restrict__ inv_scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // over n_out * (1 + k_valid)
  if (i >= n_out * (1 + k_valid)) return;
  const int n = i / (1 + k_valid), j = i - n * (1 + k_valid);  // j == 0: bias, else k = j - 1
  const int col = j == 0 ? 0 : 8 + (j - 1);
  const float* p = partial + ((int64_t)(n / 64) * 64 + (n % 64)) * NC + col;
  float s = 0.f;
  for (int x = 0; x < gx; ++x) s += __ldg(p + (int64_t)x * gy * 64 * NC);
  if (inv_scale) s *= __ldg(inv_scale);
  if (j == 0) {
    if (db) db[n] = s;
  } else if (dW) {
    dW[(int64_t)n * ldw + (j - 1)] = s;
  }
}

// ---- element-wise pieces around the GEMMs ------------------------------------------------------------
// Chunk-major fp16 features [prefix (n_prefix fp32 columns) | frequency encoding of xyz or dir]; the frequency block is
// models/encodings.py:437-444: [x (if include_input), per dim j: cos(x_j 2^k) k<n, sin(x_j 2^k) k<n]; zero padded.
__global__ void __launch_bounds__(256) k_pe_h(PointSrc src, int64_t P, int which, int n_freq, int incl, const float* __restrict__ prefix,
                                              int64_t prefix_ld, int n_prefix, __half* __restrict__ out, int64_t ld, int chunk0, int nchunks) {
  const int width = n_prefix + (incl ? 3 : 0) + 6 * n_freq;
  const int64_t total = P * nchunks;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i / P);
    const int64_t p = i - (int64_t)c * P;
    float x[3];
    if (which == 0)
      load_xyz(src, p, x[0], x[1], x[2]);
    else
      load_dir(src, p, x[0], x[1], x[2]);
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      int f = c * 8 + e;
      float val = 0.f;
      if (f < n_prefix) {
        val = __ldg(prefix + p * prefix_ld + f);
      } else if (f < width) {
        int g = f - n_prefix;
        if (incl && g < 3) {
          val = x[g];
        } else {
          g -= incl ? 3 : 0;
          int j = g / (2 * n_freq), rr = g - j * 2 * n_freq, k = rr < n_freq ? rr : rr - n_freq;
          float a = x[j] * exp2f((float)k);
          val = rr < n_freq ? cosf(a) : sinf(a);
        }
      }
      v[e] = fminf(fmaxf(val, -65504.f), 65504.f);
    }
    uint4 o = make_uint4(pack_h2(v[0], v[1]), pack_h2(v[2], v[3]), pack_h2(v[4], v[5]), pack_h2(v[6], v[7]));
    *reinterpret_cast<uint4*>(out + ((int64_t)(chunk0 + c) * ld + p) * 8) = o;
  }
}

__device__ __forceinline__ float pe_trunc_exp(float v) { return expf(fminf(fmaxf(v, -88.722839111f), 88.722839111f)); }

// out (P,4) = [rgb_act(rgb_raw), trunc_exp(sigma_raw)]  (meta_vanilla.py:109-154, trunc_exp.py:30-57)
__global__ void __launch_bounds__(256) k_pe_final_fwd(const float* __restrict__ rgb_raw, int64_t rgb_ld, const float* __restrict__ heads,
                                                      int64_t heads_ld, int sigma_col, int use_sigmoid, int64_t P, float* __restrict__ out) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float r0 = rgb_raw[p * rgb_ld], r1 = rgb_raw[p * rgb_ld + 1], r2 = rgb_raw[p * rgb_ld + 2];
    if (use_sigmoid) {
      r0 = 1.f / (1.f + expf(-r0));
      r1 = 1.f / (1.f + expf(-r1));
      r2 = 1.f / (1.f + expf(-r2));
    }
    *reinterpret_cast<float4*>(out + p * 4) = make_float4(r0, r1, r2, pe_trunc_exp(heads[p * heads_ld + sigma_col]));
  }
}

// d_out (P,4) -> scaled fp16 gradient of the colour output layer (2 chunks: 3 features + zeros), scaled d sigma_raw (P,),
// and scale[0] = s, scale[1] = 1/s with s the power of two that puts max|d_out| near 2^10.
__global__ void __launch_bounds__(256) k_pe_final_bwd(const float* __restrict__ d_out, const float* __restrict__ out, int use_sigmoid, int64_t P,
                                                      const float* __restrict__ gmax, __half* __restrict__ g_rgb, int64_t ld,
                                                      float* __restrict__ dsraw, float* __restrict__ scale) {
  float s = 1.f;
  {
    float gm = __ldg(gmax);
    if (gm > 0.f && gm < 3.0e38f) s = exp2f(fminf(fmaxf(floorf(log2f(1024.f / gm)), -100.f), 100.f));
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    scale[0] = s;
    scale[1] = 1.f / s;
  }
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x) {
    float4 dy = __ldg(reinterpret_cast<const float4*>(d_out + p * 4));
    float4 y = __ldg(reinterpret_cast<const float4*>(out + p * 4));
    float g0 = dy.x, g1 = dy.y, g2 = dy.z;
    if (use_sigmoid) {
      g0 *= y.x * (1.f - y.x);
      g1 *= y.y * (1.f - y.y);
      g2 *= y.z * (1.f - y.z);
    }
    g0 = fminf(fmaxf(g0 * s, -65504.f), 65504.f);
    g1 = fminf(fmaxf(g1 * s, -65504.f), 65504.f);
    g2 = fminf(fmaxf(g2 * s, -65504.f), 65504.f);
    *reinterpret_cast<uint4*>(g_rgb + ((int64_t)0 * ld + p) * 8) = make_uint4(pack_h2(g0, g1), pack_h2(g2, 0.f), 0u, 0u);
    *reinterpret_cast<uint4*>(g_rgb + ((int64_t)1 * ld + p) * 8) = make_uint4(0u, 0u, 0u, 0u);
    dsraw[p] = dy.w * y.w * s;  // trunc_exp backward: g * exp(clamped x) = g * sigma
  }
}

// feature `f` of a chunk-major fp16 tensor <- fp32 vector (the head gradient's sigma column)
__global__ void __launch_bounds__(256) k_set_feature(__half* __restrict__ t, int64_t ld, int f, const float* __restrict__ v, int64_t P) {
  for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < P; p += (int64_t)gridDim.x * blockDim.x)
    t[((int64_t)(f / 8) * ld + p) * 8 + (f & 7)] = __float2half_rn(fminf(fmaxf(v[p], -65504.f), 65504.f));
}

static int dense_attrs() {
  static thread_local int done_dev = -1;
  int dev = 0;
  ACN_CUDA(cudaGetDevice(&dev));
  if (done_dev != dev) {
    ACN_CUDA(cudaFuncSetAttribute(k_dense_tc<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    ACN_CUDA(cudaFuncSetAttribute(k_dense_tc<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    ACN_CUDA(cudaFuncSetAttribute(k_wgrad_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    done_dev = dev;
  }
  return ACN_OK;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_pe_encode_h(const AcnPoints* pts, int which, int n_freq, int include_input, const float* prefix, int64_t prefix_ld, int n_prefix,
                    void* out_h, int64_t ld, int chunk0, void* stream) {
  ACN_REQUIRE(pts && out_h && (pts->x6 || (pts->rays && pts->t && pts->S >= 1)), "acn_pe_encode_h: bad points / output");
  ACN_REQUIRE(n_freq >= 0 && n_freq <= 24 && (which == 0 || which == 1) && n_prefix >= 0 && (n_prefix == 0 || prefix),
              "acn_pe_encode_h: bad encoding arguments");
  ACN_REQUIRE(ld % 128 == 0 && ld >= pts->P, "acn_pe_encode_h: ld must be a multiple of 128 covering P");
  if (pts->P == 0) return ACN_OK;
  int width = n_prefix + (include_input ? 3 : 0) + 6 * n_freq, nchunks = (width + 7) / 8;
  k_pe_h<<<persistent_grid((pts->P * nchunks + 255) / 256, 8), 256, 0, as_stream(stream)>>>(
      make_src(pts), pts->P, which, n_freq, include_input, prefix, prefix_ld, n_prefix, reinterpret_cast<__half*>(out_h), ld, chunk0, nchunks);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_dense_tc(const void* a_h, int k_chunks, int64_t M, int64_t ld, const float* W, int64_t ldw, int transposed, const float* bias, int n_out,
                 int k_valid, int act, const void* mask_h, int64_t mask_ld, int mask_chunks, void* out_h, int64_t out_ld, int out_chunk0,
                 int out_chunks, float* out_f, int64_t out_f_ld, void* stream) {
  ACN_REQUIRE(a_h && W && M >= 0 && k_chunks >= 1 && k_chunks <= 46 && n_out >= 1 && n_out <= 4096 && k_valid >= 1 && k_valid <= 8 * k_chunks,
              "acn_dense_tc: bad shapes (K <= 368)");
  ACN_REQUIRE(ld % 128 == 0 && ld >= M && (out_h || out_f) && (act == 0 || act == 1), "acn_dense_tc: bad arguments");
  ACN_REQUIRE(!out_h || (out_ld % 128 == 0 && out_ld >= M && out_chunks >= 1), "acn_dense_tc: bad fp16 output");
  ACN_REQUIRE(!mask_h || (mask_ld % 128 == 0 && mask_ld >= M && mask_chunks >= 1), "acn_dense_tc: bad mask");
  if (M == 0) return ACN_OK;
  int rc = dense_attrs();
  if (rc) return rc;
  DenseArgs a;
  a.A = reinterpret_cast<const __half*>(a_h);
  a.ld = ld;
  a.M = M;
  a.KC = k_chunks;
  a.CT = (k_chunks + 2 + 1) / 2 * 2;
  a.W = W;
  a.bias = bias;
  a.ldw = ldw;
  a.n_out = n_out;
  a.k_valid = k_valid;
  a.transposed = transposed;
  a.act = act;
  a.mask = reinterpret_cast<const __half*>(mask_h);
  a.mask_ld = mask_ld;
  a.mask_chunks = mask_chunks;
  a.out_h = reinterpret_cast<__half*>(out_h);
  a.out_ld = out_ld;
  a.out_chunk0 = out_chunk0;
  a.out_chunks = out_chunks;
  a.out_f = out_f;
  a.out_f_ld = out_f_ld;
  const int NB = n_out > 32 ? 64 : 16;
  const int gy = (n_out + NB - 1) / NB;
  int64_t pairs = ((M + 127) / 128 + 1) / 2;
  int gx = num_sms_cached() / gy;
  if (gx < 1) gx = 1;
  if (pairs < gx) gx = (int)pairs;
  size_t smem = (size_t)a.CT * NB * 16 + 2 * (size_t)a.CT * DCH;
  dim3 grid(gx, gy);
  if (NB == 64)
    k_dense_tc<64><<<grid, 256, smem, as_stream(stream)>>>(a);
  else
    k_dense_tc<16><<<grid, 256, smem, as_stream(stream)>>>(a);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int64_t acn_wgrad_tc_ws_bytes(int n_out, int k_chunks) {
  int gy = (n_out + 63) / 64;
  int gx = num_sms_cached() / gy;
  if (gx < 1) gx = 1;
  return (int64_t)sizeof(float) * gx * gy * 64 * (8 + 8 * k_chunks);
}

int acn_wgrad_tc(const void* dz_h, int dz_chunks, const void* a_h, int k_chunks, int64_t M, int64_t ld, int n_out, int k_valid, float* dW,
                 int64_t ldw, float* db, const float* inv_scale_dev, void* ws, void* stream) {
  ACN_REQUIRE(dz_h && a_h && ws && M >= 0 && k_chunks >= 1 && k_chunks <= 46 && dz_chunks >= 1 && n_out >= 1 && k_valid >= 1 &&
                  k_valid <= 8 * k_chunks && n_out <= 8 * dz_chunks,
              "acn_wgrad_tc: bad shapes");
  ACN_REQUIRE(ld % 128 == 0 && ld >= M, "acn_wgrad_tc: ld must be a multiple of 128 covering M");
  int rc = dense_attrs();
  if (rc) return rc;
  cudaStream_t st = as_stream(stream);
  int gy = (n_out + 63) / 64;
  int gx = num_sms_cached() / gy;
  if (gx < 1) gx = 1;
  int64_t tiles = (M + 127) / 128;
  if (tiles < gx) gx = tiles < 1 ? 1 : (int)tiles;
  WgradTcArgs a;
  a.dZ = reinterpret_cast<const __half*>(dz_h);
  a.A = reinterpret_cast<const __half*>(a_h);
  a.ld = ld;
  a.M = M;
  a.KC = k_chunks;
  a.dz_chunks = dz_chunks;
  a.partial = reinterpret_cast<float*>(ws);
  size_t smem = (size_t)(8 + 1 + k_chunks) * DCH;
  k_wgrad_tc<<<dim3(gx, gy), 256, smem, st>>>(a);
  ACN_LAUNCH_CHECK();
  int n = n_out * (1 + k_valid);
  k_wgrad_tc_reduce<<<(n + 255) / 256, 256, 0, st>>>(a.partial, gx, gy, 8 + 8 * k_chunks, n_out, k_valid, dW, ldw, db, inv_scale_dev);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_pe_final_fwd(const float* rgb_raw, int64_t rgb_ld, const float* heads, int64_t heads_ld, int sigma_col, int use_sigmoid, int64_t P,
                     float* out, void* stream) {
  ACN_REQUIRE(rgb_raw && heads && out && P >= 0 && rgb_ld >= 3 && sigma_col >= 0 && sigma_col < heads_ld, "acn_pe_final_fwd: bad arguments");
  if (P == 0) return ACN_OK;
  k_pe_final_fwd<<<persistent_grid((P + 255) / 256, 8), 256, 0, as_stream(stream)>>>(rgb_raw, rgb_ld, heads, heads_ld, sigma_col, use_sigmoid, P, out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_pe_final_bwd(const float* d_out, const float* out, int use_sigmoid, int64_t P, void* g_rgb_h, int64_t ld, float* dsraw, float* scale2,
                     void* stream) {
  ACN_REQUIRE(d_out && out && g_rgb_h && dsraw && scale2 && P >= 0 && ld % 128 == 0 && ld >= P, "acn_pe_final_bwd: bad arguments");
  if (P == 0) return ACN_OK;
  cudaStream_t st = as_stream(stream);
  float* gmax = scale2 + 2;  // scale2 holds 4 floats: s, 1/s, max|d_out|, spare
  ACN_CUDA(cudaMemsetAsync(gmax, 0, sizeof(float), st));
  int rc = launch_absmax(d_out, P, gmax, st);
  if (rc) return rc;
  k_pe_final_bwd<<<persistent_grid((P + 255) / 256, 8), 256, 0, st>>>(d_out, out, use_sigmoid, P, gmax, reinterpret_cast<__half*>(g_rgb_h), ld,
                                                                      dsraw, scale2);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_set_feature_h(void* t_h, int64_t ld, int feature, const float* v, int64_t P, void* stream) {
  ACN_REQUIRE(t_h && v && feature >= 0 && P >= 0 && ld % 128 == 0 && ld >= P, "acn_set_feature_h: bad arguments");
  if (P == 0) return ACN_OK;
  k_set_feature<<<persistent_grid((P + 255) / 256, 8), 256, 0, as_stream(stream)>>>(reinterpret_cast<__half*>(t_h), ld, feature, v, P);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
