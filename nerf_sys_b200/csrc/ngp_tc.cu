// a11-a14 on the 5th-generation tensor cores: the hash-grid expert MLPs (MetaNGP.density / .color,
// models/inr/meta_ngp.py:171-241) as fused tcgen05 kernels.  fp16 operands (activations, weights),
// fp32 accumulation in TMEM; gradient tiles are fp16 too, multiplied by a per-launch power-of-two scale derived
// from max|d_out| (k_absmax) so they sit high in the fp16 range, saturated at +-65504, and unscaled exactly on the way out.
// This is the precision class the reference itself runs on a GPU (TF32 matmuls, nerf_runner.py:42-43,
// fp16 under autocast, meta_core.py:38); the fp32 CUDA-core kernels in ngp.cu stay as the "exact" mode.
//
//   k_ngp_mlp_fwd_tc : 5 independent 128-row pipelines per CTA; per layer one thread issues the MMAs
//                      (A = activations in smem, B = weights in smem, D in TMEM), the pipeline's 128 threads
//                      read their row back with tcgen05.ld, apply ReLU / trunc_exp / sigmoid and write the
//                      next layer's A operand in place.  Biases ride along as an extra K chunk ([1,0..] x [b,0..]).
//   k_ngp_mlp_bwd_tc : recomputes the forward for a 128-row tile (no activation round trip through HBM),
//                      then runs the dgrad chain; every weight gradient dW = dZ^T A is an M=64 MMA over the
//                      SAME smem buffers read MN-major, accumulated in TMEM across all tiles of the CTA and
//                      written once as a per-CTA partial; k_tc_wgrad_reduce sums partials in a fixed order.
// CTAs are partitioned over experts in proportion to their tile counts, so a CTA loads one expert's
// weights once and the partial sums of an expert live in a contiguous CTA range (deterministic reduce).
#include "hashgrid.cuh"
#include "tc.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace acn {

#ifndef ACN_TC_PROFILE
#define ACN_TC_PROFILE 0
#endif
#ifndef ACN_TC_G_BF16
#define ACN_TC_G_BF16 0  // mixed fp16 x bf16 operands in one kind::f16 MMA raise an illegal-instruction fault on sm_100a (measured)
#endif

using namespace tc;

constexpr uint32_t ROWS = 128;
constexpr uint32_t CH = ROWS * 16;  // bytes of one 8-column chunk of a 128-row activation tile

// smem weight block: B operands, chunk-major; chunk 0 = [bias,0,..], chunks 1..in/8 = W, last chunk = 0
constexpr uint32_t OW0 = 0;                     // 64 rows x (8 + 32 + 8)
constexpr uint32_t OW1 = OW0 + 64 * 16 * 6;     // 64 rows x (8 + 64 + 8)
constexpr uint32_t OWH = OW1 + 64 * 16 * 10;    // 16 rows x (8 + 64 + 8): rows 0..14 geo_head, row 15 sigma_head
constexpr uint32_t OW3 = OWH + 16 * 16 * 10;    // 64 rows x (8 + 32 + 8): 31 inputs padded to 32
constexpr uint32_t OW4 = OW3 + 64 * 16 * 6;     // 64 rows x (8 + 64 + 8)
constexpr uint32_t OW5 = OW4 + 64 * 16 * 10;    // 16 rows x (8 + 64 + 8): rows 0..2 rgb
constexpr uint32_t W_BYTES = OW5 + 16 * 16 * 10;

struct ExpertTableTc {
  AcnNgpExpert ex[ACN_MAX_EXPERTS];
  int E;
};

// padded layer views of the nn.Linear tensors (layer ids of acnerf.h: 0 trunk.0, 1 trunk.1, 2 sigma_head,
// 3 geo_head, 4 color.0, 5 color.1, 6 color.2)
__device__ __forceinline__ float wval(const AcnNgpExpert& ex, int L, int o, int i) {
  switch (L) {
    case 0: return __ldg(ex.w[0] + o * 32 + i);
    case 1: return __ldg(ex.w[1] + o * 64 + i);
    case 2: return o < 15 ? __ldg(ex.w[3] + o * 64 + i) : __ldg(ex.w[2] + i);
    case 3: return i < 31 ? __ldg(ex.w[4] + o * 31 + i) : 0.f;
    case 4: return __ldg(ex.w[5] + o * 64 + i);
    default: return o < 3 ? __ldg(ex.w[6] + o * 64 + i) : 0.f;
  }
}
__device__ __forceinline__ float bval(const AcnNgpExpert& ex, int L, int o) {
  switch (L) {
    case 0: return __ldg(ex.b[0] + o);
    case 1: return __ldg(ex.b[1] + o);
    case 2: return o < 15 ? __ldg(ex.b[3] + o) : __ldg(ex.b[2]);
    case 3: return __ldg(ex.b[4] + o);
    case 4: return __ldg(ex.b[5] + o);
    default: return o < 3 ? __ldg(ex.b[6] + o) : 0.f;
  }
}

// Layer, row count and padded width are template parameters (the layer switch and the divisions fold away) and the loop is
// unrolled four deep, so a thread has four weight loads in flight: the run-time form walked 35 dependent global loads
// per thread, 21 k cycles (11 us) at the top of every fused MLP kernel.
template <int L, int NROWS, int IN_PAD>
__device__ __forceinline__ void load_layer_tc(uint8_t* dst, const AcnNgpExpert& ex, int tid, int nthr) {
  __half* d = reinterpret_cast<__half*>(dst);
  constexpr int total = NROWS * (IN_PAD + 16);
#pragma unroll 4
  for (int e = tid; e < total; e += nthr) {
    int c = e / (NROWS * 8), rem = e - c * NROWS * 8;
    int o = rem >> 3, j = rem & 7;
    int k = c * 8 + j - 8;
    float v;
    if (c == 0)
      v = (j == 0) ? bval(ex, L, o) : 0.f;
    else
      v = (k < IN_PAD) ? wval(ex, L, o, k) : 0.f;
    d[e] = __float2half_rn(v);
  }
}

// expert e of the parameter block, by value, through static offsets (a runtime index makes nvcc copy the whole 1.3 KB block
// to every thread's local memory)
__device__ __forceinline__ AcnNgpExpert pick_expert(const ExpertTableTc& tab, int e) {
  AcnNgpExpert ex = tab.ex[0];
#pragma unroll
  for (int k = 1; k < ACN_MAX_EXPERTS; ++k)
    if (k == e) ex = tab.ex[k];
  return ex;
}

__device__ __forceinline__ void load_weights_tc(uint8_t* sw, const AcnNgpExpert& ex, int tid, int nthr) {
  load_layer_tc<0, 64, 32>(sw + OW0, ex, tid, nthr);
  load_layer_tc<1, 64, 64>(sw + OW1, ex, tid, nthr);
  load_layer_tc<2, 16, 64>(sw + OWH, ex, tid, nthr);
  load_layer_tc<3, 64, 32>(sw + OW3, ex, tid, nthr);
  load_layer_tc<4, 64, 64>(sw + OW4, ex, tid, nthr);
  load_layer_tc<5, 16, 64>(sw + OW5, ex, tid, nthr);
}

// ---- CTA -> (expert, tile range) ------------------------------------------------------------
// Deterministic function of the segment table only (recomputed identically by the reduce kernel).
// Expert e with t_e > 0 tiles gets n_e = 1 + floor((G - nnz) * t_e / T) CTAs, the remainder goes one
// each to the first non-empty experts.  CTA j of expert e takes tiles [t_e*j/n_e, t_e*(j+1)/n_e).
struct CtaPart {
  int e;          // -1 = no work
  int first, n;   // CTA range of this expert
  int tile0, tile1;
};

// CTA ranges of every expert: first[e], n[e] (n = 0 for an empty expert); returns the number of CTAs used.
__device__ __forceinline__ int expert_ranges(const int* s_seg, int E, int G, int* first, int* n) {
  int T = (s_seg[E] - s_seg[0]) / (int)ROWS;  // s_seg may be clamped to a window of the slot list
  for (int e = 0; e < E; ++e) first[e] = n[e] = 0;
  if (T <= 0) return 0;
  int Geff = G < T ? G : T;
  int nnz = 0;
  for (int e = 0; e < E; ++e) nnz += (s_seg[e + 1] > s_seg[e]) ? 1 : 0;
  int R = Geff - nnz, given = 0;
  for (int e = 0; e < E; ++e) {
    int t = (s_seg[e + 1] - s_seg[e]) / (int)ROWS;
    n[e] = t > 0 ? 1 + (int)(((int64_t)R * t) / T) : 0;
    given += n[e];
  }
  int leftover = Geff - given, start = 0;
  for (int e = 0; e < E; ++e) {
    if (n[e] > 0 && leftover > 0) {
      ++n[e];
      --leftover;
    }
    first[e] = start;
    start += n[e];
  }
  return start;
}

__device__ __forceinline__ CtaPart cta_partition(const int* s_seg, int E, int cta, int G) {
  CtaPart r{-1, 0, 0, 0, 0};
  int first[ACN_MAX_EXPERTS], n[ACN_MAX_EXPERTS];
  expert_ranges(s_seg, E, G, first, n);
  for (int e = 0; e < E; ++e) {
    if (n[e] > 0 && cta >= first[e] && cta < first[e] + n[e]) {
      int t = (s_seg[e + 1] - s_seg[e]) / (int)ROWS, j = cta - first[e], base = s_seg[e] / (int)ROWS;
      r.e = e;
      r.first = first[e];
      r.n = n[e];
      r.tile0 = base + (int)(((int64_t)t * j) / n[e]);
      r.tile1 = base + (int)(((int64_t)t * (j + 1)) / n[e]);
      return r;
    }
  }
  return r;
}

__device__ __forceinline__ int round_up128(int64_t m) { return (int)((m + 127) / 128 * 128); }

// ---- row-wise epilogue helpers --------------------------------------------------------------
constexpr float F16_MAX = 65504.f;

// v[0..8n) -> n fp16 chunks of row `r` starting at chunk c0 of the tile at smem address `buf`
template <int NV>
__device__ __forceinline__ void store_h(uint32_t buf, uint32_t r, int c0, const float (&v)[NV]) {
#pragma unroll
  for (int c = 0; c < NV / 8; ++c)
    st_chunk(buf + (uint32_t)(c0 + c) * CH + r * 16, pack_h2_sat(v[8 * c], v[8 * c + 1]), pack_h2_sat(v[8 * c + 2], v[8 * c + 3]),
             pack_h2_sat(v[8 * c + 4], v[8 * c + 5]), pack_h2_sat(v[8 * c + 6], v[8 * c + 7]));
}
// ReLU + saturation + fp16 conversion fused into the pack (one F2FP per two elements)
template <int NV>
__device__ __forceinline__ void store_relu(uint32_t buf, uint32_t r, int c0, const float (&v)[NV]) {
#pragma unroll
  for (int c = 0; c < NV / 8; ++c)
    st_chunk(buf + (uint32_t)(c0 + c) * CH + r * 16, pack_h2_relu_sat(v[8 * c], v[8 * c + 1]), pack_h2_relu_sat(v[8 * c + 2], v[8 * c + 3]),
             pack_h2_relu_sat(v[8 * c + 4], v[8 * c + 5]), pack_h2_relu_sat(v[8 * c + 6], v[8 * c + 7]));
}
// ReLU backward: gradient chunks gated by the stored activation (> 0) of the same rows / columns
template <int NV>
__device__ __forceinline__ void store_g_gated(uint32_t buf, uint32_t r, int c0, const float (&v)[NV], uint32_t act, int a0) {
#pragma unroll
  for (int c = 0; c < NV / 8; ++c) {
    uint32_t h[4];
    ld_chunk(act + (uint32_t)(a0 + c) * CH + r * 16, h);
    st_chunk(buf + (uint32_t)(c0 + c) * CH + r * 16, h2_gate(pack_h2_sat(v[8 * c], v[8 * c + 1]), h[0]),
             h2_gate(pack_h2_sat(v[8 * c + 2], v[8 * c + 3]), h[1]), h2_gate(pack_h2_sat(v[8 * c + 4], v[8 * c + 5]), h[2]),
             h2_gate(pack_h2_sat(v[8 * c + 6], v[8 * c + 7]), h[3]));
  }
}
template <int NV>
__device__ __forceinline__ void store_g(uint32_t buf, uint32_t r, int c0, const float (&v)[NV]) {
#pragma unroll
  for (int c = 0; c < NV / 8; ++c) {
#if ACN_TC_G_BF16
    st_chunk(buf + (uint32_t)(c0 + c) * CH + r * 16, pack_bf2(v[8 * c], v[8 * c + 1]), pack_bf2(v[8 * c + 2], v[8 * c + 3]),
             pack_bf2(v[8 * c + 4], v[8 * c + 5]), pack_bf2(v[8 * c + 6], v[8 * c + 7]));
#else
    st_chunk(buf + (uint32_t)(c0 + c) * CH + r * 16, pack_h2_sat(v[8 * c], v[8 * c + 1]), pack_h2_sat(v[8 * c + 2], v[8 * c + 3]),
             pack_h2_sat(v[8 * c + 4], v[8 * c + 5]), pack_h2_sat(v[8 * c + 6], v[8 * c + 7]));
#endif
  }
}
__device__ __forceinline__ void store_zero_chunk(uint32_t buf, uint32_t r, int c) { st_chunk(buf + (uint32_t)c * CH + r * 16, 0u, 0u, 0u, 0u); }
__device__ __forceinline__ void store_ones_chunk(uint32_t buf, uint32_t r, int c) {
  st_chunk(buf + (uint32_t)c * CH + r * 16, 0x00003C00u /* fp16 (1, 0) */, 0u, 0u, 0u);
}

// ---- MMA issue helpers (called by one thread; operand addresses in 16-byte units, see smem_desc16) --------
constexpr uint32_t CH16 = CH / 16;
// forward layer: D(128 x N) = A(128 x (8+IN+8), K-major) * W(N x (8+IN+8), K-major)^T
template <int IN, int N>
__device__ __forceinline__ void issue_fwd(uint32_t d_tmem, uint32_t a16, uint32_t w16) {
  constexpr uint32_t id = idesc_f16(128, N, F16, F16, K_MAJOR, K_MAJOR);
#pragma unroll
  for (int s = 0; s < (IN + 16) / 16; ++s)
    mma(d_tmem, smem_desc16(a16 + s * 2 * CH16, CH, 128), smem_desc16(w16 + s * 2 * N, N * 16, 128), id, s > 0);
}

constexpr uint32_t GFMT = ACN_TC_G_BF16 ? BF16 : F16;

// dgrad: D(128 x IN) = G(128 x OUT, K-major) * W(OUT x IN) ; W read MN-major from the forward weight block
// (skipping its bias chunk).  WN = rows of the weight block (= padded OUT).
template <int OUT, int IN, int WN>
__device__ __forceinline__ void issue_dgrad(uint32_t d_tmem, uint32_t g16, uint32_t w16) {
  constexpr uint32_t id = idesc_f16(128, IN, GFMT, F16, K_MAJOR, MN_MAJOR);
#pragma unroll
  for (int s = 0; s < OUT / 16; ++s)
    mma(d_tmem, smem_desc16(g16 + s * 2 * CH16, CH, 128), smem_desc16(w16 + WN + s * 16, 128, WN * 16), id, s > 0);
}

// wgrad: ACC(64 x (8+IN)) += G(128 x 64)^T * [1 | A](128 x (8+IN)), both read MN-major, K = 128 rows
template <int IN>
__device__ __forceinline__ void issue_wgrad(uint32_t acc_tmem, uint32_t g16, uint32_t a16, uint32_t accumulate) {
  constexpr uint32_t id = idesc_f16(64, 8 + IN, GFMT, F16, MN_MAJOR, MN_MAJOR);
#pragma unroll
  for (int s = 0; s < 8; ++s) mma(acc_tmem, smem_desc16(g16 + s * 16, 128, CH), smem_desc16(a16 + s * 16, 128, CH), id, accumulate | (s > 0));
}

__device__ __forceinline__ float trunc_exp_f(float v) { return expf(fminf(fmaxf(v, -88.722839111f), 88.722839111f)); }
__device__ __forceinline__ float sigmoid_f(float v) { return 1.f / (1.f + expf(-v)); }

// cin = [geo 0..14 | sh 0..15 | 0] : first / second half as fp16 chunks
__device__ __forceinline__ void dir_to_sh16(float dx, float dy, float dz, float* sh) {
  normalize3(dx, dy, dz, 1e-9f);  // meta_ngp.py:165-168
  normalize3(dx, dy, dz, 1e-9f);  // encodings.py:141
  sh16_eval(dx, dy, dz, sh);
}
__device__ __forceinline__ void dir_sh16(const PointSrc& src, int64_t p, float* sh) {
  float dx = 0.f, dy = 0.f, dz = 1.f;
  if (p >= 0) load_dir(src, p, dx, dy, dz);
  normalize3(dx, dy, dz, 1e-9f);  // meta_ngp.py:165-168
  normalize3(dx, dy, dz, 1e-9f);  // encodings.py:141
  sh16_eval(dx, dy, dz, sh);
}

// =============================================================================================
// forward
// =============================================================================================
constexpr int FWD_PIPES = 7;
constexpr int FUSED_PIPES = 5, FUSED_MAXREG = 96;  // the gather needs registers for two levels of reads in flight (measured: 4 pipelines 37.3 ms per 1080p frame, 5: 35.1, 6: 35.6, 7: 37.3)
constexpr uint32_t ABUF = 10 * CH;  // chunk 0 ones | chunks 1..8 data | chunk 9 zero (chunk 5 zero for 32-wide inputs)
constexpr size_t fwd_smem(int pipes) { return W_BYTES + (size_t)pipes * ABUF; }
constexpr uint32_t fwd_tmem_cols(int pipes) { return pipes * 64 <= 256 ? 256u : 512u; }  // power of two >= 64 * pipes

// FUSED = false: the encoding comes from HBM (`enc`, written by k_ngp_hash_fwd); a CTA serves ONE expert (cta_partition).
//         The training forward: the backward needs the encoding anyway.
// FUSED = true : the hash-grid gather runs inside the kernel - the thread that owns a row of the 128-row tile walks the 16
//         levels of its sample (two levels of table reads in flight, as k_ngp_hash_fwd) and writes the 32 features straight
//         into the fp16 A-operand chunks; the (32, ld) fp32 encoding never exists (at 1080p: 2.2 GB written and 2.8 GB read
//         back per 17 M-slot block).  While one pipeline gathers, the others run their MMA / epilogue steps.  The grid walks
//         the experts ONE AFTER ANOTHER (weights reloaded per expert), so that one 128 MiB table is hot in L2 at a time
//         instead of all of them.  Inference only (no_grad rendering).
template <int FWD_PIPES, int MAXREG, bool FUSED>
__global__ void __launch_bounds__(128 * FWD_PIPES, 1) __maxnreg__(MAXREG)
    k_ngp_mlp_fwd_tc(PointSrc src, const int32_t* __restrict__ pair_point, const float* __restrict__ pair_w, const int32_t* __restrict__ seg,
                     int64_t M_ident, ExpertTableTc tab, const float* __restrict__ enc, int64_t ld, float* __restrict__ out, GridParams g) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar[FWD_PIPES];
  __shared__ uint32_t s_tmem;
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1];
  __shared__ CtaPart s_part;
  __shared__ float s_box[6];
  __shared__ const float2* s_table;
  const int tid = threadIdx.x, q = tid >> 7;
  const uint32_t r = tid & 127;
  // a pair list truncated at its speculative capacity (route_pack without the host read) is consumed up to its last whole tile
  if (tid <= tab.E) s_seg[tid] = seg ? min(seg[tid], (int)(M_ident / ROWS) * (int)ROWS) : (tid == 0 ? 0 : round_up128(M_ident));
  __syncthreads();
  const int64_t M = seg ? (int64_t)s_seg[tab.E] : M_ident;
  if (tid == 0) {
    s_part = cta_partition(s_seg, tab.E, blockIdx.x, gridDim.x);
    for (int i = 0; i < FWD_PIPES; ++i) mbar_init(smem_u32(&s_bar[i]), 1);
    mbar_init_fence();
  }
  constexpr int FWD_THREADS = 128 * FWD_PIPES;
  constexpr uint32_t FWD_TMEM_COLS = fwd_tmem_cols(FWD_PIPES);
  if (tid < 32) tmem_alloc(&s_tmem, FWD_TMEM_COLS);
  __syncthreads();
  const CtaPart part = s_part;
  const uint32_t sw = smem_u32(smem), abuf = sw + W_BYTES + q * ABUF;
  const uint32_t sw16 = sw >> 4, abuf16 = sw16 + (W_BYTES + q * ABUF) / 16;  // smem base is 128-byte aligned
  store_ones_chunk(abuf, r, 0);
  store_zero_chunk(abuf, r, 9);
  const uint32_t tbase = s_tmem;
  const uint32_t dcol = tbase + q * 64;                       // this pipeline's accumulator columns
  const uint32_t trow = dcol + ((uint32_t)((tid >> 5) & 3) << 21);  // + lane base (32 * warp-in-group) << 16
  const uint32_t bar = smem_u32(&s_bar[q]);
  uint32_t phase = 0;
  const uint32_t hmask = (1u << g.log2_T) - 1u;

  // FUSED: every expert in turn, tiles dealt round-robin over (CTA, pipeline); else: the one expert of this CTA
  const int e_first = FUSED ? 0 : part.e, e_last = FUSED ? tab.E : part.e + 1;
  for (int e = e_first; e < e_last && e >= 0; ++e) {
    int tile_begin, tile_end, tile_step;
    if (FUSED) {
      const int t0 = s_seg[e] / (int)ROWS, t1 = s_seg[e + 1] / (int)ROWS;
      if (t1 <= t0) continue;  // CTA-uniform
      tile_begin = t0 + (int)blockIdx.x * FWD_PIPES + q;
      tile_end = t1;
      tile_step = (int)gridDim.x * FWD_PIPES;
      __syncthreads();  // every pipeline is done with the previous expert's weights
      if (tid == 0) {
        const AcnNgpExpert ex = pick_expert(tab, e);
        s_table = reinterpret_cast<const float2*>(ex.table);
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          s_box[a] = ex.box_min[a];
          s_box[3 + a] = ex.extent[a];
        }
      }
    } else {
      tile_begin = part.tile0 + q;
      tile_end = part.tile1;
      tile_step = FWD_PIPES;
    }
    load_weights_tc(smem, pick_expert(tab, e), tid, FWD_THREADS);
    fence_async_smem();
    fence_before();
    __syncthreads();
    fence_after();

    for (int tile = tile_begin; tile < tile_end; tile += tile_step) {
      const int64_t m = (int64_t)tile * ROWS + r;
      const bool in_range = m < M;
      const int64_t p = in_range ? (pair_point ? (int64_t)pair_point[m] : m) : -1;
      const int64_t mm = in_range ? m : 0;
      if (FUSED) {  // gather -> chunks 1..4 (8 features = 4 levels per chunk), chunk 5 zero
        if (p >= 0) {
          const float2* __restrict__ table = s_table;
          float x, y, z;
          load_xyz(src, p, x, y, z);
          const float ux = to_unit(x, s_box[0], s_box[3]), uy = to_unit(y, s_box[1], s_box[4]), uz = to_unit(z, s_box[2], s_box[5]);
          Cell c0 = make_cell(ux, uy, uz, g.res[0], g.interp), c1 = c0;
          LevelFetch f0, f1;
          level_fetch(table, c0, hmask, 0u, f0);
          float v8[8];
#pragma unroll
          for (int l = 0; l < 16; ++l) {
            if (l + 1 < 16) {
              c1 = make_cell(ux, uy, uz, g.res[l + 1], g.interp);
              level_fetch(table, c1, hmask, (uint32_t)(l + 1) << g.log2_T, f1);
            }
            const float2 v = level_blend(c0, f0);
            v8[2 * (l & 3)] = v.x;
            v8[2 * (l & 3) + 1] = v.y;
            if ((l & 3) == 3) store_h<8>(abuf, r, 1 + (l >> 2), v8);
            c0 = c1;
            f0 = f1;
          }
        } else {
#pragma unroll
          for (int c = 1; c <= 4; ++c) store_zero_chunk(abuf, r, c);
        }
        store_zero_chunk(abuf, r, 5);
      } else {  // encoding -> chunks 1..4, chunk 5 zero
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) x[i] = in_range ? __ldg(enc + (int64_t)i * ld + mm) : 0.f;
        store_h<32>(abuf, r, 1, x);
        store_zero_chunk(abuf, r, 5);
      }
      fence_async_smem();
      bar_sync(1 + q, 128);
      if (r == 0) {
        fence_after();
        issue_fwd<32, 64>(dcol, abuf16, sw16 + OW0 / 16);
        commit(bar);
      }
      mbar_wait(bar, phase);
      phase ^= 1;
      fence_after();
      // two 64-wide ReLU layers share this epilogue; `next` issues the following MMAs
      float v[32];
#pragma unroll 1
      for (int layer = 0; layer < 2; ++layer) {
        tmem_ld32(trow, v);
        tmem_ld_wait();
        store_relu<32>(abuf, r, 1, v);
        tmem_ld32(trow + 32, v);
        tmem_ld_wait();
        store_relu<32>(abuf, r, 5, v);
        fence_before();
        fence_async_smem();
        bar_sync(1 + q, 128);
        if (r == 0) {
          fence_after();
          if (layer == 0)
            issue_fwd<64, 64>(dcol, abuf16, sw16 + OW1 / 16);
          else
            issue_fwd<64, 16>(dcol, abuf16, sw16 + OWH / 16);
          commit(bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        fence_after();
      }
      float sigma;
      {  // heads -> cin = [geo15 | sh16 | 0] in chunks 1..4, chunk 5 zero
        float h[16], sh[16], c[32];
        tmem_ld16(trow, h);
        tmem_ld_wait();
        sigma = trunc_exp_f(h[15]);
        dir_sh16(src, p, sh);  // (requesting the direction at the top of the tile measured 6 us slower here: 7 pipelines hide it)
#pragma unroll
        for (int i = 0; i < 15; ++i) c[i] = h[i];
#pragma unroll
        for (int i = 0; i < 16; ++i) c[15 + i] = sh[i];
        c[31] = 0.f;
        store_h<32>(abuf, r, 1, c);
        store_zero_chunk(abuf, r, 5);
      }
      fence_before();
      fence_async_smem();
      bar_sync(1 + q, 128);
      if (r == 0) {
        fence_after();
        issue_fwd<32, 64>(dcol, abuf16, sw16 + OW3 / 16);
        commit(bar);
      }
      mbar_wait(bar, phase);
      phase ^= 1;
      fence_after();
#pragma unroll 1
      for (int layer = 0; layer < 2; ++layer) {
        tmem_ld32(trow, v);
        tmem_ld_wait();
        store_relu<32>(abuf, r, 1, v);
        tmem_ld32(trow + 32, v);
        tmem_ld_wait();
        store_relu<32>(abuf, r, 5, v);
        fence_before();
        fence_async_smem();
        bar_sync(1 + q, 128);
        if (r == 0) {
          fence_after();
          if (layer == 0)
            issue_fwd<64, 64>(dcol, abuf16, sw16 + OW4 / 16);
          else
            issue_fwd<64, 16>(dcol, abuf16, sw16 + OW5 / 16);
          commit(bar);
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        fence_after();
      }
      float o8[8];
      tmem_ld8(trow, o8);
      tmem_ld_wait();
      fence_before();  // the next tile's first MMA overwrites these columns after the bar_sync above it
      float r0 = sigmoid_f(o8[0]), r1 = sigmoid_f(o8[1]), r2 = sigmoid_f(o8[2]);
      if (p >= 0) {
        if (pair_point && pair_w) {
          float w = pair_w[m];
          float* o = out + p * 4;
          atomicAdd(o + 0, r0 * w);
          atomicAdd(o + 1, r1 * w);
          atomicAdd(o + 2, r2 * w);
          atomicAdd(o + 3, sigma * w);
        } else {
          *reinterpret_cast<float4*>(out + p * 4) = make_float4(r0, r1, r2, sigma);
        }
      }
    }
  }
  fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tbase, FWD_TMEM_COLS);
}

// =============================================================================================
// backward (forward recompute + dgrad chain + TMEM-resident weight gradients)
// =============================================================================================
constexpr uint32_t B_ENC = W_BYTES;            // [1 | enc32 | 0]   6 chunks
constexpr uint32_t B_H0 = B_ENC + 6 * CH;      // [1 | h0 64 | 0]  10 chunks
constexpr uint32_t B_H1 = B_H0 + 10 * CH;
constexpr uint32_t B_CIN = B_H1 + 10 * CH;     // [1 | geo15 sh16 0 | 0]  6 chunks
constexpr uint32_t B_C0 = B_CIN + 6 * CH;
constexpr uint32_t B_C1 = B_C0 + 10 * CH;
constexpr uint32_t B_GA = B_C1 + 10 * CH;      // gradient tiles, 8 chunks each (bf16)
constexpr uint32_t B_GB = B_GA + 8 * CH;
constexpr size_t BWD_SMEM = B_GB + 8 * CH;
// TMEM columns: D (64) then the six weight-gradient accumulators (M = 64 rows -> lanes 16i..16i+15 of each quarter)
constexpr uint32_t T_D = 0, T_A0 = 64, T_A1 = T_A0 + 40, T_AH = T_A1 + 72, T_A3 = T_AH + 72, T_A4 = T_A3 + 40, T_A5 = T_A4 + 72;
static_assert(T_A5 + 72 <= 512, "TMEM budget");
// per-CTA partial layout (floats): row-major (rows, 8+IN) blocks; column 0 = bias gradient, columns 8.. = dW
constexpr int P_A0 = 0, P_A1 = P_A0 + 64 * 40, P_AH = P_A1 + 64 * 72, P_A3 = P_AH + 16 * 72, P_A4 = P_A3 + 64 * 40, P_A5 = P_A4 + 64 * 72;
constexpr int P_FLOATS = P_A5 + 16 * 72;

// dump rows [0, 16*nq) of an M=64 accumulator (N columns) to `dst` (row-major, N floats per row).  Rows 16w..16w+15 sit in
// lanes 32w..32w+15, which the warps w, w+4, w+8, .. may all read: they split the 8-column chunks between them.
template <int N>
__device__ __forceinline__ void dump_acc(uint32_t tbase, uint32_t col, int nq, float* __restrict__ dst, int tid, int nparts, bool valid, float inv_s) {
  const int w = (tid >> 5) & 3, part = tid >> 7, l = tid & 31;
  if (part >= nparts || w >= nq) return;  // warp-uniform
  float v[8];
#pragma unroll 1
  for (int c = 8 * part; c < N; c += 8 * nparts) {
    tmem_ld8(tbase + col + c + ((uint32_t)w << 21), v);
    tmem_ld_wait();
    if (l < 16) {
      float4* d = reinterpret_cast<float4*>(dst + (16 * w + l) * N + c);
      d[0] = valid ? make_float4(v[0] * inv_s, v[1] * inv_s, v[2] * inv_s, v[3] * inv_s) : make_float4(0.f, 0.f, 0.f, 0.f);
      d[1] = valid ? make_float4(v[4] * inv_s, v[5] * inv_s, v[6] * inv_s, v[7] * inv_s) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// NQ column groups of 128 threads each: NQ = 4 (512 threads, 16 columns per thread) when the kernel has the SM to itself,
// NQ = 2 (256 threads, 2 x 16 columns per thread, half the register file) when the hash-grid scatter of the previous
// chunk runs beside it (acn_ngp_bwd_tc_pipelined).  [tile_lo, tile_hi) restricts the launch to a window of 128-slot tiles.
// Warp specialisation: 128 * NQ epilogue threads + ONE issuer warp (the last).  Measured with clock64 stamps before the
// split: the thread that issued the MMAs spent 8.7 k of a tile's 15 k cycles in issue code while the other 511 threads
// waited for it at the step barrier.  Now the epilogue warps bar.arrive on a named barrier when their operands are in
// shared memory and go straight to the accumulator mbarrier; the issuer warp bar.syncs on it, issues the step's MMAs,
// commits, and issues the weight-gradient MMAs of the layer just left while the epilogue of this step runs.
template <int NQ>
__global__ void __launch_bounds__(128 * NQ + 32, NQ == 2 ? 2 : 1)  // NQ = 2: at most half the register file
    k_ngp_mlp_bwd_tc(PointSrc src, const int32_t* __restrict__ pair_point, const float* __restrict__ pair_w, const int32_t* __restrict__ seg,
                     int64_t M_ident, ExpertTableTc tab, const float* __restrict__ enc, int64_t ld, const float* __restrict__ d_out,
                     float* __restrict__ d_enc, float* __restrict__ partial, const float* __restrict__ gmax, int tile_lo, int tile_hi,
                     long long* dbg, int dbg_tid) {
  // ACN_TC_PROFILE (compile-time, off in the product build): clock64 stamps of one thread of CTA 3 on its sixth tile, four
  // per chain step, plus whole-kernel phase marks; this is how the issue-bound steps and the slow prologue were found.
#if ACN_TC_PROFILE
  int dbg_i = 0; bool dbg_on = false;
  if (dbg && threadIdx.x == 0 && blockIdx.x == 3) dbg[100] = clock64();
#define DBG_STAMP() do { if (dbg_on && dbg_i < 120) dbg[dbg_i++] = clock64(); } while (0)
#define DBG_MARK(i) do { if (dbg && threadIdx.x == 0 && blockIdx.x == 3) dbg[i] = clock64(); } while (0)
#else
  (void)dbg; (void)dbg_tid;
#define DBG_STAMP() ((void)0)
#define DBG_MARK(i) ((void)0)
#endif
  constexpr int PASSES = 4 / NQ;  // 16-column passes per thread in a 64-column epilogue
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1];
  __shared__ CtaPart s_part;
  const int tid = threadIdx.x, q = tid >> 7;
  const uint32_t r = tid & 127;
  if (tid <= tab.E) {
    int v = seg ? seg[tid] : (tid == 0 ? 0 : round_up128(M_ident));
    s_seg[tid] = min(max(v, tile_lo * (int)ROWS), tile_hi * (int)ROWS);  // this launch's window of the slot list
  }
  __syncthreads();
  const int64_t M = seg ? (int64_t)__ldg(seg + tab.E) : M_ident;
  if (tid == 0) {
    s_part = cta_partition(s_seg, tab.E, blockIdx.x, gridDim.x);
    mbar_init(smem_u32(&s_bar), 1);
    mbar_init_fence();
  }
  __syncwarp();
  if (tid < 32) tmem_alloc(&s_tmem, 512);
  __syncthreads();
  const CtaPart part = s_part;
  const uint32_t sb = smem_u32(smem);
  constexpr int NEPI = 128 * NQ;          // epilogue threads
  const bool issuer = tid >= NEPI;        // warp NEPI/32
  if (part.e >= 0) load_weights_tc(smem, pick_expert(tab, part.e), tid, NEPI + 32);
  if (!issuer) {
    constexpr uint32_t bufs[6] = {B_ENC, B_H0, B_H1, B_CIN, B_C0, B_C1};
    constexpr int zc[6] = {5, 9, 9, 5, 9, 9};
#pragma unroll
    for (int b = 0; b < 12; ++b) {
      if (b % NQ != q) continue;
      if (b < 6)
        store_ones_chunk(sb + bufs[b], r, 0);
      else
        store_zero_chunk(sb + bufs[b - 6], r, zc[b - 6]);
    }
  }
  // gradient tiles start finite (rows 16.. of the 16-wide stages are read by the M=64 wgrad MMAs and ignored)
  for (int c = q; c < 8 && !issuer; c += NQ) {
    store_zero_chunk(sb + B_GA, r, c);
    store_zero_chunk(sb + B_GB, r, c);
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  DBG_MARK(101);
  const uint32_t tbase = s_tmem;
  const uint32_t tD = tbase + T_D;
  const uint32_t trow0 = tD + ((uint32_t)((tid >> 5) & 3) << 21);  // column 0 of this thread's row
  const uint32_t trow = trow0 + 16 * PASSES * q;                   // this thread's columns of D
  const uint32_t bar = smem_u32(&s_bar);
  uint32_t phase = 0;
  uint32_t acc = 0;  // 0 on the CTA's first tile: the weight-gradient accumulators are overwritten
  // gradient scale: the largest |d_out| lands near 2^10 in fp16 (power of two: unscaling is exact)
  float gscale = 1.f;
  {
    float gm = __ldg(gmax);
    if (gm > 0.f && gm < 3.0e38f) gscale = exp2f(fminf(fmaxf(floorf(log2f(1024.f / gm)), -100.f), 100.f));
  }
  const float inv_s = 1.f / gscale;

  // ---- issuer warp: the whole MMA program of a tile, one step per named-barrier round ----------------------------
  // The weight-gradient MMAs of a layer are issued AFTER the commit of the step that consumes the same gradient tile:
  // they run on the tensor pipe under that step's epilogue and are covered by the NEXT step's commit before their
  // operands are overwritten.  The last step's go before its commit (the next tile's prologue overwrites B_ENC).
  if (part.e >= 0 && issuer) {
    const uint32_t s16 = sb >> 4;  // the dynamic smem base is 128-byte aligned
    uint32_t iacc = 0;             // 0 on the CTA's first tile: the weight-gradient accumulators are overwritten
#define ISSUE_STEP(ISSUE, LATE)     \
  do {                              \
    bar_sync(1, NEPI + 32);         \
    if (tid == NEPI) {              \
      fence_after();                \
      ISSUE;                        \
      commit(bar);                  \
      LATE;                         \
    }                               \
    __syncwarp();                   \
  } while (0)
#define A16(B) (s16 + (B) / 16)
#pragma unroll 1
    for (int tile = part.tile0; tile < part.tile1; ++tile) {
      ISSUE_STEP((issue_fwd<32, 64>(tD, A16(B_ENC), A16(OW0))), (void)0);
      ISSUE_STEP((issue_fwd<64, 64>(tD, A16(B_H0), A16(OW1))), (void)0);
      ISSUE_STEP((issue_fwd<64, 16>(tD, A16(B_H1), A16(OWH))), (void)0);
      ISSUE_STEP((issue_fwd<32, 64>(tD, A16(B_CIN), A16(OW3))), (void)0);
      ISSUE_STEP((issue_fwd<64, 64>(tD, A16(B_C0), A16(OW4))), (void)0);
      ISSUE_STEP((issue_fwd<64, 16>(tD, A16(B_C1), A16(OW5))), (void)0);
      ISSUE_STEP((issue_dgrad<16, 64, 16>(tD, A16(B_GA), A16(OW5))), (issue_wgrad<64>(tbase + T_A5, A16(B_GA), A16(B_C1), iacc)));
      ISSUE_STEP((issue_dgrad<64, 64, 64>(tD, A16(B_GB), A16(OW4))), (issue_wgrad<64>(tbase + T_A4, A16(B_GB), A16(B_C0), iacc)));
      ISSUE_STEP((issue_dgrad<64, 32, 64>(tD, A16(B_GA), A16(OW3))), (issue_wgrad<32>(tbase + T_A3, A16(B_GA), A16(B_CIN), iacc)));
      ISSUE_STEP((issue_dgrad<16, 64, 16>(tD, A16(B_GB), A16(OWH))), (issue_wgrad<64>(tbase + T_AH, A16(B_GB), A16(B_H1), iacc)));
      ISSUE_STEP((issue_dgrad<64, 64, 64>(tD, A16(B_GA), A16(OW1))), (issue_wgrad<64>(tbase + T_A1, A16(B_GA), A16(B_H0), iacc)));
      ISSUE_STEP((issue_dgrad<64, 32, 64>(tD, A16(B_GB), A16(OW0)), issue_wgrad<32>(tbase + T_A0, A16(B_GB), A16(B_ENC), iacc)), (void)0);
      iacc = 1;
    }
#undef A16
#undef ISSUE_STEP
  }

  // ---- epilogue warps.  One chain step: publish this thread's smem writes to the async proxy, count the warp in on the
  // issuer's barrier (no wait), do OVERLAP work while the MMAs run, wait for the accumulator.  ISSUE / LATE only name
  // what the issuer warp runs for this step (see above).
#define TC_STEP3(ISSUE, LATE, OVERLAP) \
  do {                                 \
    DBG_STAMP();                       \
    fence_before();                    \
    fence_async_smem();                \
    bar_arrive(1, NEPI + 32);          \
    DBG_STAMP();                       \
    OVERLAP;                           \
    DBG_STAMP();                       \
    mbar_wait(bar, phase);             \
    phase ^= 1;                        \
    fence_after();                     \
    DBG_STAMP();                       \
  } while (0)
#define TC_STEP(ISSUE, LATE) TC_STEP3(ISSUE, LATE, (void)0)
  // 64-column epilogues: ReLU + mask capture (forward recompute) / mask application (dgrad chain)
#define EPI_RELU(BUF)                                                   \
  do {                                                                  \
    _Pragma("unroll") for (int s_ = 0; s_ < PASSES; ++s_) {             \
      tmem_ld16(trow + 16 * s_, v);                                     \
      tmem_ld_wait();                                                   \
      store_relu<16>(sb + BUF, r, 1 + 2 * (PASSES * q + s_), v);        \
    }                                                                   \
  } while (0)
#define EPI_MASK(BUF, ACT)                                                                         \
  do {                                                                                             \
    _Pragma("unroll") for (int s_ = 0; s_ < PASSES; ++s_) {                                        \
      tmem_ld16(trow + 16 * s_, v);                                                                \
      tmem_ld_wait();                                                                              \
      store_g_gated<16>(sb + BUF, r, 2 * (PASSES * q + s_), v, sb + ACT, 1 + 2 * (PASSES * q + s_)); \
    }                                                                                              \
  } while (0)

  if (part.e >= 0 && !issuer) {
    // software prefetch: the next tile's encoding rows and pair id are requested one tile ahead, this tile's output
    // gradient at the top of the tile, so no global-load latency sits in the per-layer chain
    constexpr int FE = 8 * PASSES;  // encoding features per thread
    float xn[FE];
    int64_t pn = -1;
    auto prefetch = [&](int tile) {
      const int64_t m_ = (int64_t)tile * ROWS + r;
      const bool ok_ = tile < part.tile1 && m_ < M;
      if (ok_) {
#pragma unroll
        for (int i = 0; i < FE; ++i) xn[i] = ldg_now(enc + (int64_t)(FE * q + i) * ld + m_);
        pn = pair_point ? (int64_t)ldg_now(pair_point + m_) : m_;
      } else {
#pragma unroll
        for (int i = 0; i < FE; ++i) xn[i] = 0.f;
        pn = -1;
      }
    };
    prefetch(part.tile0);
    for (int tile = part.tile0; tile < part.tile1; ++tile) {
      const int64_t m = (int64_t)tile * ROWS + r;
      const int64_t p = pn;
#if ACN_TC_PROFILE
      dbg_on = dbg && tid == dbg_tid && blockIdx.x == 3 && tile == part.tile0 + 5;
#endif
      DBG_STAMP();
      float v[16];
      float4 dy = make_float4(0.f, 0.f, 0.f, 0.f);
      // view direction: requested here, consumed under the heads MMA three steps later (the load used to sit in that
      // step's overlap slot and held the whole CTA for ~900 cycles)
      float ddx = 0.f, ddy = 0.f, ddz = 1.f;
      if (q == 1 && p >= 0) load_dir(src, p, ddx, ddy, ddz);
      // ---- forward recompute ----
#pragma unroll
      for (int s_ = 0; s_ < PASSES; ++s_) {
        float x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = xn[8 * s_ + i];
        store_h<8>(sb + B_ENC, r, 1 + PASSES * q + s_, x);
      }
      if (q == 0 && p >= 0) {
        dy = ldg_now(reinterpret_cast<const float4*>(d_out + p * 4));
        if (pair_w) {
          float w = __ldg(pair_w + m);
          dy.x *= w; dy.y *= w; dy.z *= w; dy.w *= w;
        }
      }
      prefetch(tile + 1);
      TC_STEP((issue_fwd<32, 64>(tD, sb + B_ENC, sb + OW0)), (void)0);
      EPI_RELU(B_H0);
      TC_STEP((issue_fwd<64, 64>(tD, sb + B_H0, sb + OW1)), (void)0);
      EPI_RELU(B_H1);
      float sh[16];
      TC_STEP3((issue_fwd<64, 16>(tD, sb + B_H1, sb + OWH)), (void)0, if (q == 1) dir_to_sh16(ddx, ddy, ddz, sh));
      float sigma = 0.f;
      if (q == 0) {  // cin[0..15] = geo 0..14, sh 0
        tmem_ld16(trow0, v);
        tmem_ld_wait();
        sigma = trunc_exp_f(v[15]);
        v[15] = 0.28209479177387814f;
        store_h<16>(sb + B_CIN, r, 1, v);
      } else if (q == 1) {  // cin[16..31] = sh 1..15, 0
#pragma unroll
        for (int i = 0; i < 15; ++i) v[i] = sh[1 + i];
        v[15] = 0.f;
        store_h<16>(sb + B_CIN, r, 3, v);
      }
      TC_STEP((issue_fwd<32, 64>(tD, sb + B_CIN, sb + OW3)), (void)0);
      EPI_RELU(B_C0);
      TC_STEP((issue_fwd<64, 64>(tD, sb + B_C0, sb + OW4)), (void)0);
      EPI_RELU(B_C1);
      TC_STEP((issue_fwd<64, 16>(tD, sb + B_C1, sb + OW5)), (void)0);
      // ---- output gradients ----
      float dsraw = 0.f;
      if (q < 2) {
        float g[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) g[i] = 0.f;
        if (q == 0) {
          float o8[8];
          tmem_ld8(trow0, o8);
          tmem_ld_wait();
          if (p >= 0) {
            float r0 = sigmoid_f(o8[0]), r1 = sigmoid_f(o8[1]), r2 = sigmoid_f(o8[2]);
            g[0] = gscale * dy.x * r0 * (1.f - r0);
            g[1] = gscale * dy.y * r1 * (1.f - r1);
            g[2] = gscale * dy.z * r2 * (1.f - r2);
            dsraw = gscale * dy.w * sigma;  // trunc_exp backward: g * exp(clamped x)  (trunc_exp.py:44-57)
          }
        }
        store_g<8>(sb + B_GA, r, q, g);  // chunk 0 = dz5 (3 of 8), chunk 1 = 0
      }
      // ---- dgrad chain; the weight gradient of the layer just left is issued behind the commit ----
      TC_STEP((issue_dgrad<16, 64, 16>(tD, sb + B_GA, sb + OW5)), (issue_wgrad<64>(tbase + T_A5, sb + B_GA, sb + B_C1, acc)));
      EPI_MASK(B_GB, B_C1);
      TC_STEP((issue_dgrad<64, 64, 64>(tD, sb + B_GB, sb + OW4)), (issue_wgrad<64>(tbase + T_A4, sb + B_GB, sb + B_C0, acc)));
      EPI_MASK(B_GA, B_C0);
      TC_STEP((issue_dgrad<64, 32, 64>(tD, sb + B_GA, sb + OW3)), (issue_wgrad<32>(tbase + T_A3, sb + B_GA, sb + B_CIN, acc)));
      if (q == 0) {  // head gradient [d geo 0..14 | d sigma_raw] -> chunks 0, 1
        tmem_ld16(trow0, v);
        tmem_ld_wait();
        v[15] = dsraw;
        store_g<16>(sb + B_GB, r, 0, v);
      }
      TC_STEP((issue_dgrad<16, 64, 16>(tD, sb + B_GB, sb + OWH)), (issue_wgrad<64>(tbase + T_AH, sb + B_GB, sb + B_H1, acc)));
      EPI_MASK(B_GA, B_H1);
      TC_STEP((issue_dgrad<64, 64, 64>(tD, sb + B_GA, sb + OW1)), (issue_wgrad<64>(tbase + T_A1, sb + B_GA, sb + B_H0, acc)));
      EPI_MASK(B_GB, B_H0);
      // last step of the tile: its weight-gradient MMAs go BEFORE the commit, because the next tile's first
      // epilogue overwrites the encoding buffer they read
      TC_STEP((issue_dgrad<64, 32, 64>(tD, sb + B_GB, sb + OW0), issue_wgrad<32>(tbase + T_A0, sb + B_GB, sb + B_ENC, acc)), (void)0);
      if (q < 2) {
        tmem_ld16(trow0 + 16 * q, v);
        tmem_ld_wait();
        if (m < ld) {
#pragma unroll
          for (int i = 0; i < 16; ++i) d_enc[(int64_t)(16 * q + i) * ld + m] = v[i] * inv_s;
        }
      }
      DBG_STAMP();
      acc = 1;
    }
  }
#undef TC_STEP
#undef TC_STEP3
#undef EPI_RELU
#undef EPI_MASK
  DBG_MARK(102);
  fence_before();
  __syncthreads();
  fence_after();
  {
    float* dst = partial + (int64_t)blockIdx.x * P_FLOATS;
    const bool valid = acc != 0;
    dump_acc<40>(tbase, T_A0, 4, dst + P_A0, tid, NQ, valid, inv_s);
    dump_acc<72>(tbase, T_A1, 4, dst + P_A1, tid, NQ, valid, inv_s);
    dump_acc<72>(tbase, T_AH, 1, dst + P_AH, tid, NQ, valid, inv_s);
    dump_acc<40>(tbase, T_A3, 4, dst + P_A3, tid, NQ, valid, inv_s);
    dump_acc<72>(tbase, T_A4, 4, dst + P_A4, tid, NQ, valid, inv_s);
    dump_acc<72>(tbase, T_A5, 1, dst + P_A5, tid, NQ, valid, inv_s);
  }
  DBG_MARK(103);
  fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tbase, 512);
}

// max |x| over a tensor (non-negative floats order like their bit patterns; NaN sorts above inf -> scale 1)
__global__ void __launch_bounds__(256) k_absmax(const float4* __restrict__ x, int64_t n4, float* __restrict__ out, const int32_t* __restrict__ bound) {
  if (bound && n4 > (int64_t)__ldg(bound)) n4 = __ldg(bound);  // rows past the used part of an owner's slab are stale
  float m = 0.f;
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v = __ldg(x + i);
    m = fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fmaxf(fabsf(v.z), fabsf(v.w)), m));
    if (!(v.x == v.x && v.y == v.y && v.z == v.z && v.w == v.w)) m = __int_as_float(0x7fc00000);
  }
  int b = __float_as_int(m);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) b = max(b, __shfl_xor_sync(0xffffffffu, b, o));
  if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int*>(out), b);
}

int launch_absmax(const float* x, int64_t n4, float* out, cudaStream_t st) {
  k_absmax<<<persistent_grid((n4 + 255) / 256, 4), 256, 0, st>>>(reinterpret_cast<const float4*>(x), n4, out, nullptr);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

// ---- fixed-order reduction of the per-CTA partials into the nn.Linear-shaped gradients -------
struct GradPtrsTc {
  float* w[ACN_MAX_EXPERTS][ACN_NGP_LAYERS];
  float* b[ACN_MAX_EXPERTS][ACN_NGP_LAYERS];
};

constexpr int TC_MAX_CHUNKS = 8;

// chunk c of a pipelined backward covers tiles [c*chunk_tiles, (c+1)*chunk_tiles) (the last chunk is open-ended); its
// partials are partial[c][cta][P_FLOATS].  Chunks and CTAs are summed in a fixed order.
__global__ void __launch_bounds__(256) k_tc_wgrad_reduce(const float* __restrict__ partial, const int32_t* __restrict__ seg, int64_t M_ident, int E,
                                                         int G, int n_chunks, int chunk_tiles, const GradPtrsTc gp) {
  __shared__ int s_seg[ACN_MAX_EXPERTS + 1];
  __shared__ int s_first[TC_MAX_CHUNKS][ACN_MAX_EXPERTS], s_n[TC_MAX_CHUNKS][ACN_MAX_EXPERTS];
  if ((int)threadIdx.x < n_chunks) {
    const int c = threadIdx.x;
    const int lo = c * chunk_tiles * (int)ROWS, hi = (c == n_chunks - 1) ? 0x7fffff80 : (c + 1) * chunk_tiles * (int)ROWS;
    int sg[ACN_MAX_EXPERTS + 1];
    for (int k = 0; k <= E; ++k) {
      int v = seg ? seg[k] : (k == 0 ? 0 : round_up128(M_ident));
      sg[k] = min(max(v, lo), hi);
    }
    expert_ranges(sg, E, G, s_first[c], s_n[c]);
  }
  (void)s_seg;
  __syncthreads();
  const int e = blockIdx.y;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= P_FLOATS) return;
  float acc = 0.f;
  for (int c = 0; c < n_chunks; ++c) {
    const float* src = partial + ((int64_t)c * G + s_first[c][e]) * P_FLOATS + j;
    // eight loads in flight, added in CTA order (the sum order, hence the result, is unchanged; the serial form was
    // one L2 round trip per partial: ~25 us for 148 CTAs)
    const int n = s_n[c][e];
    int x = 0;
    for (; x + 8 <= n; x += 8) {
      float t[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) t[u] = __ldg(src + (int64_t)(x + u) * P_FLOATS);
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += t[u];
    }
    for (; x < n; ++x) acc += __ldg(src + (int64_t)x * P_FLOATS);
  }
  // j -> (block, row, col)
  int blk, row, col, N;
  if (j < P_A1) { blk = 0; N = 40; row = (j - P_A0) / N; col = (j - P_A0) % N; }
  else if (j < P_AH) { blk = 1; N = 72; row = (j - P_A1) / N; col = (j - P_A1) % N; }
  else if (j < P_A3) { blk = 2; N = 72; row = (j - P_AH) / N; col = (j - P_AH) % N; }
  else if (j < P_A4) { blk = 3; N = 40; row = (j - P_A3) / N; col = (j - P_A3) % N; }
  else if (j < P_A5) { blk = 4; N = 72; row = (j - P_A4) / N; col = (j - P_A4) % N; }
  else { blk = 5; N = 72; row = (j - P_A5) / N; col = (j - P_A5) % N; }
  if (col >= 1 && col < 8) return;
  const bool is_b = col == 0;
  const int i = col - 8;
  int layer, o = row, in_real;
  switch (blk) {
    case 0: layer = 0; in_real = 32; break;
    case 1: layer = 1; in_real = 64; break;
    case 2: if (row < 15) { layer = 3; } else { layer = 2; o = 0; } in_real = 64; break;
    case 3: layer = 4; in_real = 31; break;
    case 4: layer = 5; in_real = 64; break;
    default: if (row >= 3) return; layer = 6; in_real = 64; break;
  }
  if (is_b) {
    float* d = gp.b[e][layer];
    if (d) d[o] = acc;
  } else if (i < in_real) {
    float* d = gp.w[e][layer];
    if (d) d[o * in_real + i] = acc;
  }
}

// ---- descriptor self-test: the three operand-layout combinations the kernels above rely on ----
//   D1 = X Y^T        X (128,64) fp16 K-major, Y (64,64) fp16 K-major                  (forward)
//   D3 = G Y          G (128,64) K-major, Y read MN-major                              (dgrad)
//   D2 = G^T [X]      G and X read MN-major, M = 64 accumulator                        (wgrad)
// G is bf16 when `mixed` (mixed fp16 x bf16 operands), fp16 otherwise.
__global__ void __launch_bounds__(128, 1) k_tc_selftest(const float* __restrict__ X, const float* __restrict__ Y, const float* __restrict__ Gm, int mixed,
                                                        float* __restrict__ D1, float* __restrict__ D2, float* __restrict__ D3) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  const int tid = threadIdx.x;
  const uint32_t r = tid;
  const uint32_t sb = smem_u32(smem), bx = sb, by = sb + 8 * CH, bg = by + 8 * 64 * 16;
  if (tid == 0) {
    mbar_init(smem_u32(&s_bar), 1);
    mbar_init_fence();
  }
  __syncwarp();
  if (tid < 32) tmem_alloc(&s_tmem, 256);
  {
    float v[32];
    for (int h = 0; h < 2; ++h) {
      for (int i = 0; i < 32; ++i) v[i] = X[r * 64 + 32 * h + i];
      store_h<32>(bx, r, 4 * h, v);
      for (int i = 0; i < 32; ++i) v[i] = Gm[r * 64 + 32 * h + i];
      if (mixed) {
        for (int c = 0; c < 4; ++c)
          st_chunk(bg + (uint32_t)(4 * h + c) * CH + r * 16, pack_bf2(v[8 * c], v[8 * c + 1]), pack_bf2(v[8 * c + 2], v[8 * c + 3]),
                   pack_bf2(v[8 * c + 4], v[8 * c + 5]), pack_bf2(v[8 * c + 6], v[8 * c + 7]));
      } else {
        store_h<32>(bg, r, 4 * h, v);
      }
    }
    if (r < 64) {  // Y: 64 rows, chunk stride 64*16
      for (int c = 0; c < 8; ++c) {
        for (int i = 0; i < 8; ++i) v[i] = Y[r * 64 + 8 * c + i];
        st_chunk(by + (uint32_t)c * (64 * 16) + r * 16, pack_h2(v[0], v[1]), pack_h2(v[2], v[3]), pack_h2(v[4], v[5]), pack_h2(v[6], v[7]));
      }
    }
  }
  fence_async_smem();
  fence_before();
  __syncthreads();
  fence_after();
  const uint32_t tbase = s_tmem;
  const uint32_t gf = mixed ? BF16 : F16;
  if (tid == 0) {
    const uint32_t id1 = idesc_f16(128, 64, F16, F16, K_MAJOR, K_MAJOR);
    for (int s = 0; s < 4; ++s) mma(tbase + 0, smem_desc(bx + s * 2 * CH, CH, 128), smem_desc(by + s * 2 * (64 * 16), 64 * 16, 128), id1, s > 0);
    const uint32_t id3 = idesc_f16(128, 64, gf, F16, K_MAJOR, MN_MAJOR);
    for (int s = 0; s < 4; ++s) mma(tbase + 64, smem_desc(bg + s * 2 * CH, CH, 128), smem_desc(by + s * 256, 128, 64 * 16), id3, s > 0);
    const uint32_t id2 = idesc_f16(64, 64, gf, F16, MN_MAJOR, MN_MAJOR);
    for (int s = 0; s < 8; ++s) mma(tbase + 128, smem_desc(bg + s * 256, 128, CH), smem_desc(bx + s * 256, 128, CH), id2, s > 0);
    commit(smem_u32(&s_bar));
  }
  mbar_wait(smem_u32(&s_bar), 0);
  fence_after();
  const uint32_t lane_off = (uint32_t)(tid >> 5) << 21;
  float v[32];
  for (int h = 0; h < 2; ++h) {
    tmem_ld32(tbase + 0 + 32 * h + lane_off, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) D1[r * 64 + 32 * h + i] = v[i];
    tmem_ld32(tbase + 64 + 32 * h + lane_off, v);
    tmem_ld_wait();
    for (int i = 0; i < 32; ++i) D3[r * 64 + 32 * h + i] = v[i];
    tmem_ld32(tbase + 128 + 32 * h + lane_off, v);
    tmem_ld_wait();
    if ((tid & 31) < 16) {
      int o = 16 * (tid >> 5) + (tid & 31);
      for (int i = 0; i < 32; ++i) D2[o * 64 + 32 * h + i] = v[i];
    }
  }
  fence_before();
  __syncthreads();
  if (tid < 32) tmem_dealloc(tbase, 256);
}

static void fill_table_tc(ExpertTableTc& tab, const AcnNgpExpert* experts, int E) {
  tab.E = E;
  for (int e = 0; e < E; ++e) tab.ex[e] = experts[e];
  for (int e = E; e < ACN_MAX_EXPERTS; ++e) tab.ex[e] = experts[0];
}

static int tc_grid(int64_t Mcap) {
  int64_t tiles = (Mcap + ROWS - 1) / ROWS;
  int g = num_sms_cached();
  if (tiles < g) g = (int)tiles;
  return g < 1 ? 1 : g;
}

static int set_tc_attrs() {
  static thread_local int done_dev = -1;
  int dev = 0;
  ACN_CUDA(cudaGetDevice(&dev));
  if (done_dev != dev) {
    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_fwd_tc<FWD_PIPES, 72, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd_smem(FWD_PIPES)));
    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_fwd_tc<FUSED_PIPES, FUSED_MAXREG, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)fwd_smem(FUSED_PIPES)));

    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_bwd_tc<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BWD_SMEM));
    ACN_CUDA(cudaFuncSetAttribute(k_ngp_mlp_bwd_tc<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BWD_SMEM));
    ACN_CUDA(cudaFuncSetAttribute(k_tc_selftest, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    done_dev = dev;
  }
  return ACN_OK;
}

}  // namespace acn

using namespace acn;

extern "C" {

int acn_ngp_mlp_fwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const float* enc, float* out,
                       void* stream) {
  ACN_REQUIRE(pts && pairs && experts && enc && out, "acn_ngp_mlp_fwd_tc: NULL argument");
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "acn_ngp_mlp_fwd_tc: E must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->S >= 1), "acn_ngp_mlp_fwd_tc: points need x6 or rays");
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? (E == 1 && pairs->Mcap == pts->P) : (pairs->seg != nullptr), "acn_ngp_mlp_fwd_tc: inconsistent pair list");
  ACN_REQUIRE(pairs->Mcap < ((int64_t)1 << 31) - 128, "acn_ngp_mlp_fwd_tc: chunk the pair list below 2^31");
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  int rc = set_tc_attrs();
  if (rc) return rc;
  ExpertTableTc tab;
  fill_table_tc(tab, experts, E);
  int64_t ld = (Mcap + 127) / 128 * 128;
  GridParams gp{};
  k_ngp_mlp_fwd_tc<FWD_PIPES, 72, false><<<tc_grid(Mcap), 128 * FWD_PIPES, fwd_smem(FWD_PIPES), as_stream(stream)>>>(
      make_src(pts), pairs->point, pairs->weight, pairs->seg, Mcap, tab, enc, ld, out, gp);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int acn_ngp_fused_fwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g, float* out,
                         void* stream) {
  ACN_REQUIRE(pts && pairs && experts && g && out, "acn_ngp_fused_fwd_tc: NULL argument");
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "acn_ngp_fused_fwd_tc: E must be in [1,%d]", ACN_MAX_EXPERTS);
  ACN_REQUIRE(g->levels == 16 && g->features == 2 && g->log2_T >= 1 && g->log2_T <= 24 && g->interp >= 0 && g->interp <= 1,
              "acn_ngp_fused_fwd_tc: needs L=16, F=2, linear or smoothstep interpolation");
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->t && pts->S >= 1), "acn_ngp_fused_fwd_tc: points need x6 or (rays,t,S)");
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? (E == 1 && pairs->Mcap == pts->P) : (pairs->seg != nullptr), "acn_ngp_fused_fwd_tc: inconsistent pair list");
  ACN_REQUIRE(pairs->Mcap < ((int64_t)1 << 31) - 128, "acn_ngp_fused_fwd_tc: chunk the pair list below 2^31");
  for (int e = 0; e < E; ++e) ACN_REQUIRE(experts[e].table, "acn_ngp_fused_fwd_tc: expert %d has no table", e);
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  int rc = set_tc_attrs();
  if (rc) return rc;
  ExpertTableTc tab;
  fill_table_tc(tab, experts, E);
  const PointSrc src = make_src(pts);
  const GridParams gp = make_grid(g);
  cudaStream_t st = as_stream(stream);
  k_ngp_mlp_fwd_tc<FUSED_PIPES, FUSED_MAXREG, true><<<tc_grid(Mcap), 128 * FUSED_PIPES, fwd_smem(FUSED_PIPES), st>>>(
      src, pairs->point, pairs->weight, pairs->seg, Mcap, tab, nullptr, 0, out, gp);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

int64_t acn_ngp_bwd_tc_ws_bytes(void) {
  return (int64_t)sizeof(float) * (P_FLOATS * (int64_t)num_sms_cached() + 4);
}

static int run_bwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g, const float* enc,
                      const float* d_out, const AcnNgpExpertGrads* grads, float* d_enc, void* ws, int n_chunks, cudaStream_t st,
                      cudaStream_t side, const char* who) {
  ACN_REQUIRE(pts && pairs && experts && enc && d_out && grads && d_enc && ws, "%s: NULL argument", who);
  ACN_REQUIRE(E >= 1 && E <= ACN_MAX_EXPERTS, "%s: E must be in [1,%d]", who, ACN_MAX_EXPERTS);
  ACN_REQUIRE(pts->x6 || (pts->rays && pts->S >= 1), "%s: points need x6 or rays", who);
  bool ident = pairs->point == nullptr;
  ACN_REQUIRE(ident ? (E == 1 && pairs->Mcap == pts->P) : (pairs->seg != nullptr), "%s: inconsistent pair list", who);
  ACN_REQUIRE(pairs->Mcap < ((int64_t)1 << 31) - 128, "%s: chunk the pair list below 2^31", who);
  ACN_REQUIRE(n_chunks >= 1 && n_chunks <= TC_MAX_CHUNKS, "%s: 1..%d chunks", who, TC_MAX_CHUNKS);
  int64_t Mcap = pairs->Mcap;
  if (Mcap == 0) return ACN_OK;
  int rc = set_tc_attrs();
  if (rc) return rc;
  ExpertTableTc tab;
  fill_table_tc(tab, experts, E);
  int64_t ld = (Mcap + 127) / 128 * 128;
  int G = tc_grid(Mcap);
  float* partial = reinterpret_cast<float*>(ws);
  GradPtrsTc gp;
  bool any = false;
  for (int e = 0; e < ACN_MAX_EXPERTS; ++e)
    for (int l = 0; l < ACN_NGP_LAYERS; ++l) {
      gp.w[e][l] = e < E ? grads[e].w[l] : nullptr;
      gp.b[e][l] = e < E ? grads[e].b[l] : nullptr;
      any = any || gp.w[e][l] || gp.b[e][l];
    }
  float* gmax = partial + (int64_t)P_FLOATS * G * n_chunks;
  ACN_CUDA(cudaMemsetAsync(gmax, 0, sizeof(float), st));
  const bool store_mode = pairs->point && !pairs->weight;  // point id == row of d_out: rows at or past seg[E] are unused
  k_absmax<<<persistent_grid((pts->P + 255) / 256, 4), 256, 0, st>>>(reinterpret_cast<const float4*>(d_out), pts->P, gmax,
                                                                     store_mode ? pairs->seg + E : nullptr);
  ACN_LAUNCH_CHECK();
  const int tiles = (int)(ld / ROWS);
  const int chunk_tiles = (tiles + n_chunks - 1) / n_chunks;
  static thread_local cudaEvent_t ev[TC_MAX_CHUNKS + 2];  // fork / per-chunk / join markers (never used for timing)
  static thread_local int ev_dev = -1;
  long long* dbgbuf = nullptr;
  int dbg_tid = 0;
#if ACN_TC_PROFILE
  static long long* s_dbg = nullptr;
  if (getenv("ACN_TC_DBG")) {
    if (!s_dbg) { cudaMalloc(&s_dbg, 128 * 8); cudaMemset(s_dbg, 0, 128 * 8); }
    dbgbuf = s_dbg;
    dbg_tid = atoi(getenv("ACN_TC_DBG"));
  }
#endif
  if (!side) {  // the kernel has the SM to itself
    ACN_REQUIRE(n_chunks == 1, "%s: chunks need a side stream", who);
    k_ngp_mlp_bwd_tc<4><<<G, 512 + 32, BWD_SMEM, st>>>(make_src(pts), pairs->point, pairs->weight, pairs->seg, Mcap, tab, enc, ld, d_out, d_enc,
                                                  partial, gmax, 0, tiles, dbgbuf, dbg_tid);
    ACN_LAUNCH_CHECK();
#if ACN_TC_PROFILE
    if (dbgbuf) {
      cudaStreamSynchronize(st);
      long long hb[128];
      cudaMemcpy(hb, dbgbuf, sizeof(hb), cudaMemcpyDeviceToHost);
      fprintf(stderr, "TCDBG");
      for (int i = 0; i < 50; ++i) fprintf(stderr, " %lld", hb[i] - hb[100]);
      fprintf(stderr, "\nTCKERN setup %lld loop %lld dump %lld\n", hb[101] - hb[100], hb[102] - hb[101], hb[103] - hb[102]);
    }
#endif
  } else {
    // Software pipeline over chunks of the slot list: while chunk c's table scatter (bound by the L2 atomic units, it keeps
    // its rate with 4 CTAs per SM) runs on the side stream, chunk c+1's MLP backward (bound by its tcgen05 layer chain, half
    // the register file) runs beside it on the same SMs.
    int dev = 0;
    ACN_CUDA(cudaGetDevice(&dev));
    if (ev_dev != dev) {
      for (int i = 0; i < TC_MAX_CHUNKS + 2; ++i) ACN_CUDA(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
      ev_dev = dev;
    }
    ACN_CUDA(cudaEventRecord(ev[TC_MAX_CHUNKS], st));
    ACN_CUDA(cudaStreamWaitEvent(side, ev[TC_MAX_CHUNKS], 0));
    const int hash_gx = (num_sms_cached() * 4 + g->levels - 1) / g->levels;
    for (int c = 0; c < n_chunks; ++c) {
      const int lo = c * chunk_tiles, hi = (c == n_chunks - 1) ? tiles : (c + 1) * chunk_tiles;
      k_ngp_mlp_bwd_tc<2><<<G, 256 + 32, BWD_SMEM, st>>>(make_src(pts), pairs->point, pairs->weight, pairs->seg, Mcap, tab, enc, ld, d_out, d_enc,
                                                    partial + (int64_t)c * G * P_FLOATS, gmax, lo, hi, nullptr, 0);
      ACN_LAUNCH_CHECK();
      ACN_CUDA(cudaEventRecord(ev[c], st));
      ACN_CUDA(cudaStreamWaitEvent(side, ev[c], 0));
      rc = launch_hash_bwd_range(pts, pairs, experts, E, g, d_enc, grads, (int64_t)lo * ROWS, (int64_t)hi * ROWS, hash_gx, side);
      if (rc) return rc;
    }
    ACN_CUDA(cudaEventRecord(ev[TC_MAX_CHUNKS + 1], side));
  }
  if (any) {
    dim3 grid((P_FLOATS + 255) / 256, E);
    k_tc_wgrad_reduce<<<grid, 256, 0, st>>>(partial, pairs->seg, Mcap, E, G, n_chunks, chunk_tiles, gp);
    ACN_LAUNCH_CHECK();
  }
  if (side) ACN_CUDA(cudaStreamWaitEvent(st, ev[TC_MAX_CHUNKS + 1], 0));  // join: the scatter of the last chunk
  return ACN_OK;
}

int acn_ngp_mlp_bwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const float* enc, const float* d_out,
                       const AcnNgpExpertGrads* grads, float* d_enc, void* ws, void* stream) {
  return run_bwd_tc(pts, pairs, experts, E, nullptr, enc, d_out, grads, d_enc, ws, 1, as_stream(stream), nullptr, "acn_ngp_mlp_bwd_tc");
}

int64_t acn_ngp_bwd_tc_pipelined_ws_bytes(int n_chunks) {
  return (int64_t)sizeof(float) * (P_FLOATS * (int64_t)num_sms_cached() * n_chunks + 4);
}

int acn_ngp_bwd_tc_pipelined(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const AcnGridDesc* g,
                             const float* enc, const float* d_out, const AcnNgpExpertGrads* grads, float* d_enc, void* ws, int n_chunks,
                             void* stream, void* side_stream) {
  ACN_REQUIRE(g && side_stream && side_stream != stream, "acn_ngp_bwd_tc_pipelined: needs a grid descriptor and a distinct side stream");
  ACN_REQUIRE(g->levels == 16 && g->features == 2, "acn_ngp_bwd_tc_pipelined: fused expert path needs L=16, F=2");
  int rc = run_bwd_tc(pts, pairs, experts, E, g, enc, d_out, grads, d_enc, ws, n_chunks, as_stream(stream), as_stream(side_stream),
                      "acn_ngp_bwd_tc_pipelined");
  if (rc) return rc;
  return ACN_OK;
}

int acn_tc_selftest(const float* X, const float* Y, const float* G, int mixed, float* D1, float* D2, float* D3, void* stream) {
  ACN_REQUIRE(X && Y && G && D1 && D2 && D3, "acn_tc_selftest: NULL argument");
  int rc = set_tc_attrs();
  if (rc) return rc;
  k_tc_selftest<<<1, 128, 64 * 1024, as_stream(stream)>>>(X, Y, G, mixed, D1, D2, D3);
  ACN_LAUNCH_CHECK();
  return ACN_OK;
}

}  // extern "C"
