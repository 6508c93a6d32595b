// a8-a14 fused: hash-grid experts (MetaNGP, models/inr/meta_ngp.py:155-241) over routed
// (point, expert) pairs, forward and backward.  fp32 CUDA-core ("exact") MLP path.
//
// Pipeline per call (all stream-ordered, no host sync):
//   fwd : k_ngp_hash_fwd   pairs -> enc   (L2/HBM gather-bound, 1024 B table rows / pair)
//         k_ngp_mlp_fwd    enc,dir -> rgb,sigma ; combine out[p] (+)= w*y   (FFMA-bound, 13 440 MAC / pair)
//   bwd : k_ngp_mlp_bwd    d_out -> dz of every layer + d_enc              (FFMA-bound)
//         wgrad_launch x7  dW = dz^T x per layer, per expert               (FFMA-bound, deterministic)
//         k_ngp_hash_bwd   d_enc -> grad table scatter                     (L2 atomic-bound)
// Activations are kept feature-major [feature][slot] so thread-per-slot kernels read
// and write coalesced 128-byte rows.
#include "hashgrid.cuh"

namespace acn {

constexpr int NT = ACN_TILE;  // threads per CTA == slots per tile

// save-buffer rows (each row = ld floats)
constexpr int R_H0 = 0, R_H1 = 64, R_CIN = 128, R_C0 = 160, R_C1 = 224, R_Y = 288, R_SAVE = 292;
// backward workspace rows
constexpr int R_DZ5 = 0, R_DZ4 = 4, R_DZ3 = 68, R_DZH = 132, R_DZ1 = 148, R_DZ0 = 212, R_WS = 276;

// smem weight block (floats)
constexpr int W0_OFF = 0;                    // 64 x 32
constexpr int W1_OFF = W0_OFF + 64 * 32;     // 64 x 64
constexpr int WH_OFF = W1_OFF + 64 * 64;     // 16 x 64  (rows 0..14 geo_head, row 15 sigma_head)
constexpr int W3_OFF = WH_OFF + 16 * 64;     // 64 x 32  (31 padded to 32)
constexpr int W4_OFF = W3_OFF + 64 * 32;     // 64 x 64
constexpr int W5_OFF = W4_OFF + 64 * 64;     // 4 x 64   (3 padded to 4)
constexpr int B_OFF = W5_OFF + 4 * 64;       // biases: 64,64,16,64,64,4
constexpr int B0 = B_OFF, B1 = B0 + 64, BH = B1 + 64, B3 = BH + 16, B4 = B3 + 64, B5 = B4 + 64;
constexpr int W_FLOATS = B5 + 4;
constexpr int ACT_FLOATS = 64 * NT;
constexpr size_t MLP_SMEM = sizeof(float) * (W_FLOATS + ACT_FLOATS) + 64 * sizeof(int);

struct ExpertTable {
  AcnNgpExpert ex[ACN_MAX_EXPERTS];
  int E;
};

__device__ __forceinline__ int64_t round_ld(int64_t m) { return (m + 127) / 128 * 128; }
static inline int64_t h_round_ld(int64_t m) { return (m + 127) / 128 * 128; }

// expert owning slot m (segments are tile-aligned, so a tile never straddles experts)
__device__ __forceinline__ int find_expert(const int* s_seg, int E, int64_t m) {
  int e = 0;
  for (int k = 1; k < E; ++k)
    if (m >= s_seg[k]) e = k;
  return e;
}

// ---------------------------------------------------------------------------------------
// hash-grid forward over pairs
// ---------------------------------------------------------------------------------------
// One thread = one pair slot x LG consecutive levels.  Everything that does not depend on the level - the slot's point id,
// its expert, o + d*t, the three IEEE divisions of the unit-cube map - is computed once per LG levels (the one-level form
// spent 330 warp instructions per level, 72 % issue-slot utilisation: profiles/r2_hash_fwd_ncu.md), and the table reads of the
// LG levels are issued together, branch-free (level_fetch), before any of them is consumed.
// CTA order = expert-major, then level group, then 256-slot chunk: the CTA waves walk one expert's LG x 8 MiB slabs at a time,
// so the hot set in L2 is LG x 8 MiB instead of "one level of every expert" (64 MiB with 8 experts, which the streaming
// encoding writes kept evicting).
template <int LG>
__global__ void __launch_bounds__(256, 4) k_ngp_hash_fwd(PointSrc src, const int32_t* __restrict__ pair_point, const int32_t* __restrict__ seg,
                                                         int64_t M_ident, ExpertTable tab, GridParams g, float* __restrict__ enc, int64_t ld) {
  constexpr int NLG = 16 / LG;
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1];
  if (threadIdx.x <= tab.E) s_seg[threadIdx.x] = seg ? min(seg[threadIdx.x], (int)M_ident) : (threadIdx.x == 0 ? 0 : (int)M_ident);  // never past the capacity
  __syncthreads();
  const uint32_t mask = (1u << g.log2_T) - 1u;
  int jobs = 0;
  for (int e = 0; e < tab.E; ++e) jobs += ((s_seg[e + 1] - s_seg[e] + 255) >> 8) * NLG;
  for (int job = blockIdx.x; job < jobs; job += gridDim.x) {  // CTA-uniform
    // job -> (expert, level group, 256-slot chunk), expert-major
    int b = job, e = 0, n = 0;
    for (; e < tab.E; ++e) {
      n = (s_seg[e + 1] - s_seg[e] + 255) >> 8;
      if (b < n * NLG) break;
      b -= n * NLG;
    }
    const int lg = b / n, c = b - lg * n, l0 = lg * LG;
    const int64_t m = (int64_t)s_seg[e] + (c << 8) + threadIdx.x;
    if (m >= s_seg[e + 1]) continue;
    const int64_t p = pair_point ? (int64_t)pair_point[m] : m;
    if (p < 0) {
#pragma unroll
      for (int j = 0; j < LG; ++j) {
        __stcs(enc + (int64_t)(2 * (l0 + j)) * ld + m, 0.f);
        __stcs(enc + (int64_t)(2 * (l0 + j) + 1) * ld + m, 0.f);
      }
      continue;
    }
    const AcnNgpExpert& ex = tab.ex[e];
    const float2* __restrict__ table = reinterpret_cast<const float2*>(ex.table);
    float x, y, z;
    load_xyz(src, p, x, y, z);
    const float ux = to_unit(x, ex.box_min[0], ex.extent[0]), uy = to_unit(y, ex.box_min[1], ex.extent[1]),
                uz = to_unit(z, ex.box_min[2], ex.extent[2]);
    if (g.interp == 2) {
#pragma unroll
      for (int j = 0; j < LG; ++j) {
        Cell cc = make_cell(ux, uy, uz, g.res[l0 + j], 2);
        float2 v = level_encode(table, cc, mask, (uint32_t)(l0 + j) << g.log2_T, 2);
        __stcs(enc + (int64_t)(2 * (l0 + j)) * ld + m, v.x);
        __stcs(enc + (int64_t)(2 * (l0 + j) + 1) * ld + m, v.y);
      }
      continue;
    }
    // two levels in flight: the reads of level j+1 are issued before level j is blended
    Cell c0 = make_cell(ux, uy, uz, g.res[l0], g.interp), c1 = c0;
    LevelFetch f0, f1;
    level_fetch(table, c0, mask, (uint32_t)l0 << g.log2_T, f0);
#pragma unroll
    for (int j = 0; j < LG; ++j) {
      if (j + 1 < LG) {
        c1 = make_cell(ux, uy, uz, g.res[l0 + j + 1], g.interp);
        level_fetch(table, c1, mask, (uint32_t)(l0 + j + 1) << g.log2_T, f1);
      }
      const float2 v = level_blend(c0, f0);
      // streaming stores: the encoding is read once by the next kernel and must not push the table slabs out of L2
      __stcs(enc + (int64_t)(2 * (l0 + j)) * ld + m, v.x);
      __stcs(enc + (int64_t)(2 * (l0 + j) + 1) * ld + m, v.y);
      c0 = c1;
      f0 = f1;
    }
  }
}

struct GradTables {
  float* t[ACN_MAX_EXPERTS];
};

// Table-gradient scatter, grid.y = level (a CTA wave works on one 8 MiB slab of the gradient table, which stays L2-resident).
// Coarse levels (agg.s[l] > 0: cells wider than the spacing of consecutive samples) go through the warp-aggregated
// scatter; the reductions are REDG (explicit .global, see red_add in common.cuh).
// 4 bits per level in one 64-bit kernel parameter (a small array indexed by blockIdx.y made nvcc copy the parameter block
// to local memory in every thread: 32 KB of local stores per CTA, behind the REDs in the same LSU queue)
using AggSteps = unsigned long long;
// (the fused expert path is specialised on L = 16, see check_common)

__global__ void __launch_bounds__(256) k_ngp_hash_bwd(PointSrc src, const int32_t* __restrict__ pair_point, const int32_t* __restrict__ seg,
                                                      int64_t M_ident, ExpertTable tab, GridParams g, const float* __restrict__ denc, int64_t ld,
                                                      GradTables gts, int64_t m_lo, int64_t m_hi, AggSteps agg) {
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1];
  __shared__ float2* s_gt[ACN_MAX_EXPERTS];
  if (threadIdx.x <= tab.E) s_seg[threadIdx.x] = seg ? min(seg[threadIdx.x], (int)M_ident) : (threadIdx.x == 0 ? 0 : (int)M_ident);  // never past the capacity
#pragma unroll
  for (int k = 0; k < ACN_MAX_EXPERTS; ++k)  // static parameter offsets: a runtime index would put the block in local memory
    if (threadIdx.x == 32 + k) s_gt[k] = reinterpret_cast<float2*>(gts.t[k]);
  __syncthreads();
  int64_t M = seg ? s_seg[tab.E] : M_ident;
  const int l = blockIdx.y;
  const uint32_t mask = (1u << g.log2_T) - 1u, base = (uint32_t)l << g.log2_T;
  const int steps = (int)((agg >> (4 * l)) & 15ull);
  const uint32_t lane_key = 0xffffffffu - (threadIdx.x & 31u);
  if (m_hi < M) M = m_hi;
  for (int64_t mb = m_lo + blockIdx.x * (int64_t)blockDim.x; mb < M; mb += (int64_t)gridDim.x * blockDim.x) {  // CTA-uniform
    const int64_t m = mb + threadIdx.x;
    const bool in_range = m < M;
    const int64_t p = in_range ? (pair_point ? (int64_t)pair_point[m] : m) : -1;
    const int e = (seg && in_range) ? find_expert(s_seg, tab.E, m) : 0;
    float2* gt = s_gt[e];
    const bool live = p >= 0 && gt != nullptr;
    Cell c{0u, 0u, 0u, 0.f, 0.f, 0.f};
    float2 gv = make_float2(0.f, 0.f);
    if (live) {
      const AcnNgpExpert& ex = tab.ex[e];
      float x, y, z;
      load_xyz(src, p, x, y, z);
      c = make_cell(to_unit(x, ex.box_min[0], ex.extent[0]), to_unit(y, ex.box_min[1], ex.extent[1]),
                    to_unit(z, ex.box_min[2], ex.extent[2]), g.res[l], g.interp);
      gv = make_float2(__ldg(denc + (int64_t)(2 * l) * ld + m), __ldg(denc + (int64_t)(2 * l + 1) * ld + m));
    }
    if (steps > 0) {  // CTA-uniform; full warps
      const uint32_t key = live ? (c.ix | (c.iy << 9) | (c.iz << 18) | ((uint32_t)e << 27)) : lane_key;
      level_scatter_agg(gt, c, mask, base, gv, key, live, steps);
    } else if (live) {
      level_scatter(gt, c, mask, base, g.interp, gv);
    }
  }
}

// ---------------------------------------------------------------------------------------
// MLP helpers: thread-per-slot, weights broadcast from shared memory as float4
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ void load_weights(float* sw, const AcnNgpExpert& ex) {
  // nn.Linear (out,in) row-major tensors copied into the padded smem block
  for (int i = threadIdx.x; i < 64 * 32; i += NT) sw[W0_OFF + i] = __ldg(ex.w[0] + i);
  for (int i = threadIdx.x; i < 64 * 64; i += NT) sw[W1_OFF + i] = __ldg(ex.w[1] + i);
  for (int i = threadIdx.x; i < 15 * 64; i += NT) sw[WH_OFF + i] = __ldg(ex.w[3] + i);
  for (int i = threadIdx.x; i < 64; i += NT) sw[WH_OFF + 15 * 64 + i] = __ldg(ex.w[2] + i);
  for (int i = threadIdx.x; i < 64 * 32; i += NT) {
    int o = i >> 5, c = i & 31;
    sw[W3_OFF + i] = (c < 31) ? __ldg(ex.w[4] + o * 31 + c) : 0.f;
  }
  for (int i = threadIdx.x; i < 64 * 64; i += NT) sw[W4_OFF + i] = __ldg(ex.w[5] + i);
  for (int i = threadIdx.x; i < 4 * 64; i += NT) sw[W5_OFF + i] = (i < 3 * 64) ? __ldg(ex.w[6] + i) : 0.f;
  for (int i = threadIdx.x; i < 64; i += NT) {
    sw[B0 + i] = __ldg(ex.b[0] + i);
    sw[B1 + i] = __ldg(ex.b[1] + i);
    sw[B3 + i] = __ldg(ex.b[4] + i);
    sw[B4 + i] = __ldg(ex.b[5] + i);
    if (i < 15) sw[BH + i] = __ldg(ex.b[3] + i);
    if (i == 15) sw[BH + 15] = __ldg(ex.b[2]);
    if (i < 4) sw[B5 + i] = (i < 3) ? __ldg(ex.b[6] + i) : 0.f;
  }
}

// y[o] = b[o] + sum_i W[o][i] x[i], produced OC outputs at a time and handed to `emit(o, value)`.
template <int IN, int OUT, int OC, class Emit>
__device__ __forceinline__ void dense_fwd(const float* __restrict__ W, const float* __restrict__ b, const float (&x)[IN], Emit emit) {
  static_assert(IN % 4 == 0 && OUT % OC == 0, "shape");
#pragma unroll 1
  for (int o0 = 0; o0 < OUT; o0 += OC) {
    float acc[OC];
#pragma unroll
    for (int oo = 0; oo < OC; ++oo) acc[oo] = b[o0 + oo];
#pragma unroll
    for (int i = 0; i < IN; i += 4) {
#pragma unroll
      for (int oo = 0; oo < OC; ++oo) {
        float4 w = *reinterpret_cast<const float4*>(W + (o0 + oo) * IN + i);
        acc[oo] = fmaf(w.x, x[i], acc[oo]);
        acc[oo] = fmaf(w.y, x[i + 1], acc[oo]);
        acc[oo] = fmaf(w.z, x[i + 2], acc[oo]);
        acc[oo] = fmaf(w.w, x[i + 3], acc[oo]);
      }
    }
#pragma unroll
    for (int oo = 0; oo < OC; ++oo) emit(o0 + oo, acc[oo]);
  }
}

// dx[i] = sum_o W[o][i] dz[o], produced 8 inputs at a time.
template <int IN, int OUT, class Emit>
__device__ __forceinline__ void dense_bwd(const float* __restrict__ W, const float (&dz)[OUT], Emit emit) {
  static_assert(IN % 8 == 0, "shape");
#pragma unroll 1
  for (int i0 = 0; i0 < IN; i0 += 8) {
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
#pragma unroll
    for (int o = 0; o < OUT; ++o) {
      float4 a = *reinterpret_cast<const float4*>(W + o * IN + i0);
      float4 c = *reinterpret_cast<const float4*>(W + o * IN + i0 + 4);
      acc[0] = fmaf(a.x, dz[o], acc[0]);
      acc[1] = fmaf(a.y, dz[o], acc[1]);
      acc[2] = fmaf(a.z, dz[o], acc[2]);
      acc[3] = fmaf(a.w, dz[o], acc[3]);
      acc[4] = fmaf(c.x, dz[o], acc[4]);
      acc[5] = fmaf(c.y, dz[o], acc[5]);
      acc[6] = fmaf(c.z, dz[o], acc[6]);
      acc[7] = fmaf(c.w, dz[o], acc[7]);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) emit(i0 + k, acc[k]);
  }
}

struct TileIter {
  int64_t tile, tile_end;
};

// contiguous tile ranges per CTA so a CTA reloads weights only when the expert changes
__device__ __forceinline__ TileIter my_tiles(int64_t ntiles) {
  TileIter it;
  it.tile = ntiles * blockIdx.x / gridDim.x;
  it.tile_end = ntiles * (blockIdx.x + 1) / gridDim.x;
  return it;
}

// ---------------------------------------------------------------------------------------
// MLP forward
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT) k_ngp_mlp_fwd(PointSrc src, const int32_t* __restrict__ pair_point, const float* __restrict__ pair_w,
                                                    const int32_t* __restrict__ seg, int64_t M_ident, ExpertTable tab,
                                                    const float* __restrict__ enc, float* __restrict__ sv, int64_t ld, int keep,
                                                    float* __restrict__ out) {
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;
  float* act = smem + W_FLOATS;
  int* s_seg = reinterpret_cast<int*>(act + ACT_FLOATS);
  if (threadIdx.x <= tab.E) s_seg[threadIdx.x] = seg ? min(seg[threadIdx.x], (int)M_ident) : (threadIdx.x == 0 ? 0 : (int)M_ident);  // never past the capacity
  __syncthreads();
  int64_t M = seg ? s_seg[tab.E] : M_ident;
  TileIter it = my_tiles((M + NT - 1) / NT);
  int cur_e = -1;
  float* a = act + threadIdx.x;  // this thread's activation column: a[o*NT]
  for (; it.tile < it.tile_end; ++it.tile) {
    int64_t m = it.tile * NT + threadIdx.x;
    int e = seg ? find_expert(s_seg, tab.E, it.tile * NT) : 0;
    if (e != cur_e) {
      __syncthreads();
      load_weights(sw, tab.ex[e]);
      __syncthreads();
      cur_e = e;
    }
    bool in_range = m < M;
    int64_t p = in_range ? (pair_point ? (int64_t)pair_point[m] : m) : -1;
    int64_t mm = in_range ? m : 0;  // clamped slot for loads
    float* svm = sv + mm;  // only dereferenced when `save`
    bool save = keep && in_range;

    float x32[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) x32[i] = in_range ? __ldg(enc + (int64_t)i * ld + mm) : 0.f;
    dense_fwd<32, 64, 8>(sw + W0_OFF, sw + B0, x32, [&](int o, float v) {
      v = fmaxf(v, 0.f);
      a[o * NT] = v;
      if (save) svm[(int64_t)(R_H0 + o) * ld] = v;
    });
    float x64[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) x64[i] = a[i * NT];
    dense_fwd<64, 64, 8>(sw + W1_OFF, sw + B1, x64, [&](int o, float v) {
      v = fmaxf(v, 0.f);
      a[o * NT] = v;
      if (save) svm[(int64_t)(R_H1 + o) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) x64[i] = a[i * NT];
    // heads: rows 0..14 geo, row 15 sigma
    float sigma = 0.f;
    dense_fwd<64, 16, 8>(sw + WH_OFF, sw + BH, x64, [&](int o, float v) {
      if (o == 15) {
        // trunc_exp (models/trunc_exp.py:30-57)
        sigma = expf(fminf(fmaxf(v, -88.722839111f), 88.722839111f));
        a[31 * NT] = 0.f;
        if (save) svm[(int64_t)(R_CIN + 31) * ld] = 0.f;
      } else {
        a[o * NT] = v;
        if (save) svm[(int64_t)(R_CIN + o) * ld] = v;
      }
    });
    {
      float dx = 0.f, dy = 0.f, dz = 1.f;
      if (p >= 0) load_dir(src, p, dx, dy, dz);
      normalize3(dx, dy, dz, 1e-9f);  // meta_ngp.py:165-168
      normalize3(dx, dy, dz, 1e-9f);  // encodings.py:141
      float sh[16];
      sh16_eval(dx, dy, dz, sh);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        a[(15 + i) * NT] = sh[i];
        if (save) svm[(int64_t)(R_CIN + 15 + i) * ld] = sh[i];
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) x32[i] = a[i * NT];
    dense_fwd<32, 64, 8>(sw + W3_OFF, sw + B3, x32, [&](int o, float v) {
      v = fmaxf(v, 0.f);
      a[o * NT] = v;
      if (save) svm[(int64_t)(R_C0 + o) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) x64[i] = a[i * NT];
    dense_fwd<64, 64, 8>(sw + W4_OFF, sw + B4, x64, [&](int o, float v) {
      v = fmaxf(v, 0.f);
      a[o * NT] = v;
      if (save) svm[(int64_t)(R_C1 + o) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) x64[i] = a[i * NT];
    float rgb[4];
    dense_fwd<64, 4, 4>(sw + W5_OFF, sw + B5, x64, [&](int o, float v) { rgb[o] = 1.f / (1.f + expf(-v)); });
    if (save) {
      svm[(int64_t)(R_Y + 0) * ld] = rgb[0];
      svm[(int64_t)(R_Y + 1) * ld] = rgb[1];
      svm[(int64_t)(R_Y + 2) * ld] = rgb[2];
      svm[(int64_t)(R_Y + 3) * ld] = sigma;
    }
    if (p >= 0) {
      if (pair_point && pair_w) {
        float w = pair_w[m];
        float* o = out + p * 4;
        atomicAdd(o + 0, rgb[0] * w);
        atomicAdd(o + 1, rgb[1] * w);
        atomicAdd(o + 2, rgb[2] * w);
        atomicAdd(o + 3, sigma * w);
      } else {
        *reinterpret_cast<float4*>(out + p * 4) = make_float4(rgb[0], rgb[1], rgb[2], sigma);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// MLP backward (dgrad chain); weight gradients are reduced afterwards by wgrad_launch
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(NT) k_ngp_mlp_bwd(const int32_t* __restrict__ pair_point, const float* __restrict__ pair_w,
                                                    const int32_t* __restrict__ seg, int64_t M_ident, ExpertTable tab,
                                                    const float* __restrict__ sv, int64_t ld, const float* __restrict__ d_out,
                                                    float* __restrict__ ws, float* __restrict__ d_enc) {
  extern __shared__ __align__(16) float smem[];
  float* sw = smem;
  float* act = smem + W_FLOATS;
  int* s_seg = reinterpret_cast<int*>(act + ACT_FLOATS);
  if (threadIdx.x <= tab.E) s_seg[threadIdx.x] = seg ? min(seg[threadIdx.x], (int)M_ident) : (threadIdx.x == 0 ? 0 : (int)M_ident);  // never past the capacity
  __syncthreads();
  int64_t M = seg ? s_seg[tab.E] : M_ident;
  int64_t Mld = round_ld(M);
  TileIter it = my_tiles(Mld / NT);
  int cur_e = -1;
  float* a = act + threadIdx.x;
  for (; it.tile < it.tile_end; ++it.tile) {
    int64_t m = it.tile * NT + threadIdx.x;  // m < ld always: every ws row is fully written (zeros for padding)
    int e = seg ? find_expert(s_seg, tab.E, it.tile * NT) : 0;
    if (e != cur_e) {
      __syncthreads();
      load_weights(sw, tab.ex[e]);
      __syncthreads();
      cur_e = e;
    }
    bool in_range = m < M;
    int64_t p = in_range ? (pair_point ? (int64_t)pair_point[m] : m) : -1;
    const float* svm = sv + (in_range ? m : 0);
    float* wsm = ws + m;
    float4 dy = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p >= 0) {
      dy = __ldg(reinterpret_cast<const float4*>(d_out + p * 4));
      if (pair_w) {
        float w = pair_w[m];
        dy.x *= w; dy.y *= w; dy.z *= w; dy.w *= w;
      }
    }
    bool live = p >= 0;
    float r0 = svm[(int64_t)(R_Y + 0) * ld], r1 = svm[(int64_t)(R_Y + 1) * ld], r2 = svm[(int64_t)(R_Y + 2) * ld];
    float sg = svm[(int64_t)(R_Y + 3) * ld];
    float dz5[4];
    dz5[0] = live ? dy.x * r0 * (1.f - r0) : 0.f;
    dz5[1] = live ? dy.y * r1 * (1.f - r1) : 0.f;
    dz5[2] = live ? dy.z * r2 * (1.f - r2) : 0.f;
    dz5[3] = 0.f;
    float dsraw = live ? dy.w * sg : 0.f;  // trunc_exp backward: g * exp(clamped)
#pragma unroll
    for (int i = 0; i < 4; ++i) wsm[(int64_t)(R_DZ5 + i) * ld] = dz5[i];

    // dz4 = (W5^T dz5) * [c1 > 0]
    dense_bwd<64, 4>(sw + W5_OFF, dz5, [&](int i, float v) {
      v = (live && svm[(int64_t)(R_C1 + i) * ld] > 0.f) ? v : 0.f;
      a[i * NT] = v;
      wsm[(int64_t)(R_DZ4 + i) * ld] = v;
    });
    float d64[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) d64[i] = a[i * NT];
    // dz3 = (W4^T dz4) * [c0 > 0]
    dense_bwd<64, 64>(sw + W4_OFF, d64, [&](int i, float v) {
      v = (live && svm[(int64_t)(R_C0 + i) * ld] > 0.f) ? v : 0.f;
      a[i * NT] = v;
      wsm[(int64_t)(R_DZ3 + i) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) d64[i] = a[i * NT];
    // dcin = W3^T dz3 ; geo part -> dzh[0..14], dzh[15] = d sigma_raw
    float dzh[16];
    dense_bwd<32, 64>(sw + W3_OFF, d64, [&](int i, float v) { a[i * NT] = live ? v : 0.f; });
#pragma unroll
    for (int i = 0; i < 15; ++i) dzh[i] = a[i * NT];
    dzh[15] = dsraw;
#pragma unroll
    for (int i = 0; i < 16; ++i) wsm[(int64_t)(R_DZH + i) * ld] = dzh[i];
    // dz1 = (Wh^T dzh) * [h1 > 0]
    dense_bwd<64, 16>(sw + WH_OFF, dzh, [&](int i, float v) {
      v = (live && svm[(int64_t)(R_H1 + i) * ld] > 0.f) ? v : 0.f;
      a[i * NT] = v;
      wsm[(int64_t)(R_DZ1 + i) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) d64[i] = a[i * NT];
    // dz0 = (W1^T dz1) * [h0 > 0]
    dense_bwd<64, 64>(sw + W1_OFF, d64, [&](int i, float v) {
      v = (live && svm[(int64_t)(R_H0 + i) * ld] > 0.f) ? v : 0.f;
      a[i * NT] = v;
      wsm[(int64_t)(R_DZ0 + i) * ld] = v;
    });
#pragma unroll
    for (int i = 0; i < 64; ++i) d64[i] = a[i * NT];
    // d_enc = W0^T dz0
    dense_bwd<32, 64>(sw + W0_OFF, d64, [&](int i, float v) { d_enc[(int64_t)i * ld + m] = live ? v : 0.f; });
  }
}

static int fill_table(ExpertTable& tab, const AcnNgpExpert* experts, int E) {
  tab.E = E;
  for (int e = 0; e < E; ++e) tab.ex[e] = experts[e];
  for (int e = E; e < ACN_MAX_EXPERTS; ++e) tab.ex[e] = experts[0];
  return 0;
}

static int check_common(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                        const char* who) {
  ACN_REQUIRE(pts && pairs && experts && g, "%s: NULL argument", who);
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "%s: E must be in [1,%d]", who, ACN_MAX_EXPERTS);
  ACN_REQUIRE(g->levels == 16 && g->features == 2, "%s: fused expert path needs L=16, F=2 (got L=%d F=%d)", who, g->levels, g->features);
  ACN_REQUIRE(g->log2_T >= 1 && g->log2_T <= 24 && g->interp >= 0 && g->interp <= 2, "%s: bad grid descriptor", who);
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->t && pts->S >= 1), "%s: points need x6 or (rays,t,S)", who);
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? (E == 1 && pairs->Mcap == pts->P) : (pairs->seg != nullptr), "%s: inconsistent pair list", who);
  ACN_REQUIRE(pairs->Mcap < ((int64_t)1 << 31), "%s: chunk the pair list below 2^31", who);
  return ACN_OK;
}

// table-gradient scatter over slots [m_lo, m_hi) (grid_x = 0: the default persistent grid)
int launch_hash_bwd_range(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                          const float* d_enc, const AcnNgpExpertGrads* grads, int64_t m_lo, int64_t m_hi, int grid_x, cudaStream_t st) {
  int rc = check_common(pts, pairs, experts, E, g, "acn_ngp_encode_bwd");
  if (rc) return rc;
  ACN_REQUIRE(d_enc && grads, "acn_ngp_encode_bwd: NULL argument");
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0 || m_lo >= Mcap) return ACN_OK;
  ExpertTable tab;
  fill_table(tab, experts, E);
  bool any_table = false;
  GradTables gts;
  for (int e = 0; e < ACN_MAX_EXPERTS; ++e) {
    gts.t[e] = (e < E) ? grads[e].table : nullptr;
    any_table = any_table || gts.t[e];
  }
  if (!any_table) return ACN_OK;
  int64_t span = (m_hi < Mcap ? m_hi : Mcap) - m_lo;
  // aggregation depth per level: cells of a coarse level are wider than the spacing of consecutive samples, so runs of
  // 2..8 lanes share a cell (measured at C2: 661 -> 494 us from REDG alone, -> ~410 us with the aggregation)
  AggSteps agg = 0;
  for (int l = 0; l < g->levels && l < 16; ++l) {
    int r = g->res[l];
    unsigned long long st_l = (g->interp != 2 && r < 512) ? (r <= 20 ? 3 : r <= 40 ? 2 : r <= 80 ? 1 : 0) : 0;
    agg |= st_l << (4 * l);
  }
  // (a level-group form like the gather's was measured slower: 1020 vs 777 us over 8 tables, 430 vs 340 us over one - the
  // reductions want ONE level's slabs hot in L2 at a time; profiles/r2_gather_scatter_levels.txt)
  dim3 hg(grid_x > 0 ? grid_x : persistent_grid((span + 255) / 256, 4), g->levels);
  k_ngp_hash_bwd<<<hg, 256, 0, st>>>(make_src(pts), pairs->point, pairs->seg, Mcap, tab, make_grid(g), d_enc, h_round_ld(Mcap), gts, m_lo, m_hi,
                                     agg);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // namespace acn

using namespace acn;

extern "C" {

int64_t acn_ngp_ld(int64_t Mcap) { return h_round_ld(Mcap); }
int64_t acn_ngp_save_bytes(int64_t Mcap) { return sizeof(float) * R_SAVE * h_round_ld(Mcap); }
int64_t acn_ngp_bwd_ws_bytes(int64_t Mcap, int E) { return sizeof(float) * (R_WS * h_round_ld(Mcap) + wgrad_partial_floats(E)); }

int acn_ngp_encode_fwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                       float* enc, void* stream) {
  int rc = check_common(pts, pairs, experts, E, g, "acn_ngp_encode_fwd");
  if (rc) return rc;
  ACN_REQUIRE(enc, "acn_ngp_encode_fwd: enc is NULL");
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  ExpertTable tab;
  fill_table(tab, experts, E);
  // one CTA per (expert, level group, 256-slot chunk); an expert's last chunk may be partial, hence the + E
  const int64_t chunks = (Mcap + 255) / 256 + E;
  int lg = launch_hint(ACN_HINT_GATHER_LEVELS);
  const PointSrc src = make_src(pts);
  const GridParams gp = make_grid(g);
  const int64_t ld = h_round_ld(Mcap);
  cudaStream_t st = as_stream(stream);
  const int64_t total = chunks * 16;
  const int per_sm = 4;  // 4 CTAs of 256 threads x 64 registers fill an SM
  if (lg == 0) lg = Mcap >= ((int64_t)4 << 20) ? 8 : 4;  // measured: 8 for the renderer's 17 M-slot blocks, 4 for a training batch
  switch (lg) {
    case 1: k_ngp_hash_fwd<1><<<persistent_grid(total, per_sm), 256, 0, st>>>(src, pairs->point, pairs->seg, Mcap, tab, gp, enc, ld); break;
    case 2: k_ngp_hash_fwd<2><<<persistent_grid(total / 2, per_sm), 256, 0, st>>>(src, pairs->point, pairs->seg, Mcap, tab, gp, enc, ld); break;
    case 8: k_ngp_hash_fwd<8><<<persistent_grid(total / 8, per_sm), 256, 0, st>>>(src, pairs->point, pairs->seg, Mcap, tab, gp, enc, ld); break;
    case 16: k_ngp_hash_fwd<16><<<persistent_grid(total / 16, per_sm), 256, 0, st>>>(src, pairs->point, pairs->seg, Mcap, tab, gp, enc, ld); break;
    default: k_ngp_hash_fwd<4><<<persistent_grid(total / 4, per_sm), 256, 0, st>>>(src, pairs->point, pairs->seg, Mcap, tab, gp, enc, ld); break;
  }
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_ngp_encode_bwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                       const float* d_enc, const AcnNgpExpertGrads* grads, void* stream) {
  return launch_hash_bwd_range(pts, pairs, experts, E, g, d_enc, grads, 0, (int64_t)1 << 62, 0, as_stream(stream));
}

static int set_mlp_attrs() {
  static thread_local int done_dev = -1;
  int dev = 0;
  ACN_CUDA(cudaGetDevice(&dev));
  if (done_dev != dev) {
    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM));
    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MLP_SMEM));
    done_dev = dev;
  }
  return ACN_OK;
}

int acn_ngp_mlp_fwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const float* enc, float* out,
                    void* save, void* stream) {
  ACN_REQUIRE(pts && pairs && experts && enc && out, "acn_ngp_mlp_fwd: NULL argument");
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "acn_ngp_mlp_fwd: E must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->S >= 1), "acn_ngp_mlp_fwd: points need x6 or rays");
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? (E == 1 && pairs->Mcap == pts->P) : (pairs->seg != nullptr), "acn_ngp_mlp_fwd: inconsistent pair list");
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  int rc = set_mlp_attrs();
  if (rc) return rc;
  ExpertTable tab;
  fill_table(tab, experts, E);
  int grid = persistent_grid((Mcap + NT - 1) / NT, 2);
  k_ngp_mlp_fwd<<<grid, NT, MLP_SMEM, as_stream(stream)>>>(make_src(pts), pairs->point, pairs->weight, pairs->seg, Mcap, tab, enc,
                                                            reinterpret_cast<float*>(save), h_round_ld(Mcap), save ? 1 : 0, out);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_ngp_mlp_bwd(const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const float* enc, const float* d_out, const void* save,
                    const AcnNgpExpertGrads* grads, float* d_enc, void* ws, void* stream) {
  ACN_REQUIRE(pairs && experts && enc && d_out && save && grads && d_enc && ws, "acn_ngp_mlp_bwd: NULL argument");
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "acn_ngp_mlp_bwd: E must be in [1,%d]", ACN_MAX_EXPERTS);
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? E == 1 : (pairs->seg != nullptr), "acn_ngp_mlp_bwd: inconsistent pair list");
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  int rc = set_mlp_attrs();
  if (rc) return rc;
  cudaStream_t st = as_stream(stream);
  ExpertTable tab;
  fill_table(tab, experts, E);
  int64_t ld = h_round_ld(Mcap);
  const float* sv = reinterpret_cast<const float*>(save);
  float* w = reinterpret_cast<float*>(ws);
  float* partial = w + (int64_t)R_WS * ld;
  int grid = persistent_grid(ld / NT, 2);
  k_ngp_mlp_bwd<<<grid, NT, MLP_SMEM, st>>>(pairs->point, pairs->weight, pairs->seg, Mcap, tab, sv, ld, d_out, w, d_enc);
  ACN_LAUNCH_CHECK();
  // weight gradients: layer -> (dz rows, x rows, OUT, IN); x of layer 0 is the encoding itself
  struct L {
    int layer, dz_row, x_row, OUT, IN;
  };
  const L jobs[7] = {{0, R_DZ0, -1, 64, 32},  {1, R_DZ1, R_H0, 64, 64}, {3, R_DZH, R_H1, 15, 64}, {2, R_DZH + 15, R_H1, 1, 64},
                     {4, R_DZ3, R_CIN, 64, 31}, {5, R_DZ4, R_C0, 64, 64}, {6, R_DZ5, R_C1, 3, 64}};
  for (const L& j : jobs) {
    float* dW[ACN_MAX_EXPERTS];
    float* db[ACN_MAX_EXPERTS];
    bool any = false;
    for (int e = 0; e < E; ++e) {
      dW[e] = grads[e].w[j.layer];
      db[e] = grads[e].b[j.layer];
      any = any || dW[e] || db[e];
    }
    if (!any) continue;
    WgradJob job{w + (int64_t)j.dz_row * ld, j.x_row < 0 ? enc : sv + (int64_t)j.x_row * ld, j.OUT, j.IN, ld};
    rc = wgrad_launch(job, pairs->seg, E, Mcap, dW, db, partial, st);
    if (rc) return rc;
  }
  return ACN_OK;
}

}  // extern "C"
