// Multi-resolution hash grid, torch-backend semantics of the reference
// (models/encodings.py:308-381): every level hashed, res_l = floor(min_res*g^l),
// s = x01*res, corners floor(s) and floor(s)+1 (no wrap), idx = (ix ^ iy*2654435761 ^ iz*805459861) mod T.
#pragma once
#include "common.cuh"

namespace acn {

constexpr uint32_t HASH_PY = 2654435761u;  // encodings.py:273
constexpr uint32_t HASH_PZ = 805459861u;

struct GridParams {
  int L, log2_T, interp;
  int res[ACN_MAX_LEVELS];
};

static inline GridParams make_grid(const AcnGridDesc* g) {
  GridParams p;
  p.L = g->levels;
  p.log2_T = g->log2_T;
  p.interp = g->interp;
  for (int i = 0; i < ACN_MAX_LEVELS; ++i) p.res[i] = g->res[i];
  return p;
}

// The reference hashes in int64; the low log2_T (<= 24) bits of the products and
// XOR are the same in uint32 arithmetic for the non-negative corner coordinates.
__device__ __forceinline__ uint32_t hash3(uint32_t ix, uint32_t iy, uint32_t iz, uint32_t mask) {
  return (ix ^ (iy * HASH_PY) ^ (iz * HASH_PZ)) & mask;
}

struct Cell {
  uint32_t ix, iy, iz;  // floor corner
  float wx, wy, wz;     // fractional position (after the optional smoothstep)
};

// encodings.py:334-352.  Separate multiply / floor / subtract (no FMA): the floor
// decides table rows, so it must round exactly as torch does.
__device__ __forceinline__ Cell make_cell(float x, float y, float z, int res, int interp) {
  float r = (float)res;
  float sx = __fmul_rn(x, r), sy = __fmul_rn(y, r), sz = __fmul_rn(z, r);
  Cell c;
  if (interp == 2) {  // Nearest: round half to even, as torch.round
    c.ix = (uint32_t)(int)rintf(sx);
    c.iy = (uint32_t)(int)rintf(sy);
    c.iz = (uint32_t)(int)rintf(sz);
    c.wx = c.wy = c.wz = 0.f;
    return c;
  }
  float fx = floorf(sx), fy = floorf(sy), fz = floorf(sz);
  c.ix = (uint32_t)(int)fx;
  c.iy = (uint32_t)(int)fy;
  c.iz = (uint32_t)(int)fz;
  c.wx = __fsub_rn(sx, fx);
  c.wy = __fsub_rn(sy, fy);
  c.wz = __fsub_rn(sz, fz);
  if (interp == 1) {  // Smoothstep
    c.wx = c.wx * c.wx * (3.f - 2.f * c.wx);
    c.wy = c.wy * c.wy * (3.f - 2.f * c.wy);
    c.wz = c.wz * c.wz * (3.f - 2.f * c.wz);
  }
  return c;
}

// corner c = 4*bx + 2*by + bz
__device__ __forceinline__ void corner_rows(const Cell& c, uint32_t mask, uint32_t base, uint32_t* rows) {
  uint32_t hy0 = c.iy * HASH_PY, hy1 = (c.iy + 1u) * HASH_PY;
  uint32_t hz0 = c.iz * HASH_PZ, hz1 = (c.iz + 1u) * HASH_PZ;
  uint32_t x0 = c.ix, x1 = c.ix + 1u;
  rows[0] = ((x0 ^ hy0 ^ hz0) & mask) + base;
  rows[1] = ((x0 ^ hy0 ^ hz1) & mask) + base;
  rows[2] = ((x0 ^ hy1 ^ hz0) & mask) + base;
  rows[3] = ((x0 ^ hy1 ^ hz1) & mask) + base;
  rows[4] = ((x1 ^ hy0 ^ hz0) & mask) + base;
  rows[5] = ((x1 ^ hy0 ^ hz1) & mask) + base;
  rows[6] = ((x1 ^ hy1 ^ hz0) & mask) + base;
  rows[7] = ((x1 ^ hy1 ^ hz1) & mask) + base;
}

// Gather + trilinear interpolation in the reference's x -> y -> z order (encodings.py:373-379).
__device__ __forceinline__ float2 level_encode(const float2* __restrict__ table, const Cell& c, uint32_t mask, uint32_t base,
                                               int interp) {
  if (interp == 2) return __ldg(table + (hash3(c.ix, c.iy, c.iz, mask) + base));
  uint32_t rows[8];
  corner_rows(c, mask, base, rows);
  float2 f[8];
  if ((c.ix & 1u) == 0u) {
    // even ix: the x-neighbour's row is rows[i] ^ 1 (the x prime is 1, so only bit 0 of the hash changes): both
    // corners sit in one aligned 16-byte pair -> 4 LDG.128 instead of 8 LDG.64, half the L1 wavefronts
    const float4* t4 = reinterpret_cast<const float4*>(table);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float4 q = __ldg(t4 + (rows[i] >> 1));
      bool odd = rows[i] & 1u;
      f[i] = odd ? make_float2(q.z, q.w) : make_float2(q.x, q.y);
      f[i + 4] = odd ? make_float2(q.x, q.y) : make_float2(q.z, q.w);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = __ldg(table + rows[i]);
  }
  float ax = 1.f - c.wx, ay = 1.f - c.wy, az = 1.f - c.wz;
  float2 c00, c01, c10, c11, c0, c1, o;
  c00.x = f[0].x * ax + f[4].x * c.wx;
  c00.y = f[0].y * ax + f[4].y * c.wx;
  c01.x = f[1].x * ax + f[5].x * c.wx;
  c01.y = f[1].y * ax + f[5].y * c.wx;
  c10.x = f[2].x * ax + f[6].x * c.wx;
  c10.y = f[2].y * ax + f[6].y * c.wx;
  c11.x = f[3].x * ax + f[7].x * c.wx;
  c11.y = f[3].y * ax + f[7].y * c.wx;
  c0.x = c00.x * ay + c10.x * c.wy;
  c0.y = c00.y * ay + c10.y * c.wy;
  c1.x = c01.x * ay + c11.x * c.wy;
  c1.y = c01.y * ay + c11.y * c.wy;
  o.x = c0.x * az + c1.x * c.wz;
  o.y = c0.y * az + c1.y * c.wz;
  return o;
}

// The same gather split in two, branch-free, so that a thread can have the reads of several levels in flight:
// level_fetch issues 4 LDG.128 of the aligned 16-byte pairs that hold rows[0..3] (for even ix the x-neighbour's row is the
// other half of the pair; for odd ix that half is simply unused - same sector, same wavefront as an LDG.64) plus, predicated
// on odd ix, 4 LDG.64 of rows[4..7]; level_blend picks the halves and interpolates exactly as level_encode does.
struct LevelFetch {
  float4 q[4];
  float2 h[4];
  uint32_t par;  // bit i: rows[i] is the odd half of its pair; bit 4: ix is odd
};
__device__ __forceinline__ void level_fetch(const float2* __restrict__ table, const Cell& c, uint32_t mask, uint32_t base, LevelFetch& f) {
  uint32_t rows[8];
  corner_rows(c, mask, base, rows);
  const float4* t4 = reinterpret_cast<const float4*>(table);
  const bool odd_x = (c.ix & 1u) != 0u;
  f.par = odd_x ? 16u : 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f.q[i] = __ldg(t4 + (rows[i] >> 1));
    f.par |= (rows[i] & 1u) << i;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f.h[i] = make_float2(0.f, 0.f);
    if (odd_x) f.h[i] = __ldg(table + rows[i + 4]);
  }
}
__device__ __forceinline__ float2 level_blend(const Cell& c, const LevelFetch& t) {
  float2 f[8];
  const bool odd_x = (t.par & 16u) != 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const bool odd = (t.par >> i) & 1u;
    const float2 lo = make_float2(t.q[i].x, t.q[i].y), hi = make_float2(t.q[i].z, t.q[i].w);
    f[i] = odd ? hi : lo;
    f[i + 4] = odd_x ? t.h[i] : (odd ? lo : hi);
  }
  float ax = 1.f - c.wx, ay = 1.f - c.wy, az = 1.f - c.wz;
  float2 c00, c01, c10, c11, c0, c1, o;
  c00.x = f[0].x * ax + f[4].x * c.wx;
  c00.y = f[0].y * ax + f[4].y * c.wx;
  c01.x = f[1].x * ax + f[5].x * c.wx;
  c01.y = f[1].y * ax + f[5].y * c.wx;
  c10.x = f[2].x * ax + f[6].x * c.wx;
  c10.y = f[2].y * ax + f[6].y * c.wx;
  c11.x = f[3].x * ax + f[7].x * c.wx;
  c11.y = f[3].y * ax + f[7].y * c.wx;
  c0.x = c00.x * ay + c10.x * c.wy;
  c0.y = c00.y * ay + c10.y * c.wy;
  c1.x = c01.x * ay + c11.x * c.wy;
  c1.y = c01.y * ay + c11.y * c.wy;
  o.x = c0.x * az + c1.x * c.wz;
  o.y = c0.y * az + c1.y * c.wz;
  return o;
}

// Scatter of one level's upstream gradient into the fp32 gradient table.
__device__ __forceinline__ void level_scatter(float2* __restrict__ gtable, const Cell& c, uint32_t mask, uint32_t base,
                                              int interp, float2 g) {
  if (interp == 2) {
    red_add(gtable + (hash3(c.ix, c.iy, c.iz, mask) + base), g);
    return;
  }
  uint32_t rows[8];
  corner_rows(c, mask, base, rows);
  float ax = 1.f - c.wx, ay = 1.f - c.wy, az = 1.f - c.wz;
  if ((c.ix & 1u) == 0u) {  // x-neighbours share an aligned 16-byte pair (see level_encode): 4 RED.128 instead of 8 RED.64
    float4* t4 = reinterpret_cast<float4*>(gtable);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float wyz = ((i & 2) ? c.wy : ay) * ((i & 1) ? c.wz : az);
      float w0 = ax * wyz, w1 = c.wx * wyz;
      bool odd = rows[i] & 1u;
      float4 v = odd ? make_float4(g.x * w1, g.y * w1, g.x * w0, g.y * w0) : make_float4(g.x * w0, g.y * w0, g.x * w1, g.y * w1);
      red_add(t4 + (rows[i] >> 1), v);
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float w = ((i & 4) ? c.wx : ax) * ((i & 2) ? c.wy : ay) * ((i & 1) ? c.wz : az);
    red_add(gtable + rows[i], make_float2(g.x * w, g.y * w));
  }
}

// Warp-aggregated scatter for the coarse levels.  Lanes are consecutive slots (consecutive samples of a ray), so on a
// level whose cells are wider than the sample spacing neighbouring lanes fall in the same cell.  Runs of equal `key`
// (cell + expert; a lane without work passes a key no other lane has and live = false) are cut into chunks of
// 2^steps lanes, the 8 corner x 2 feature contributions are summed inside a chunk with a bounded shuffle tree, and only
// the first lane of a chunk issues the reductions.  The RED rate (not the bytes) bounds this kernel: 215 G lane-ops/s
// on B200 whatever the operand width (profiles/r1_v5_red_microbench.txt).  Called by full warps.
__device__ __forceinline__ void level_scatter_agg(float2* __restrict__ gtable, const Cell& c, uint32_t mask, uint32_t base, float2 g,
                                                  uint32_t key, bool live, int steps) {
  const uint32_t lane = threadIdx.x & 31u;
  uint32_t prev = __shfl_up_sync(0xffffffffu, key, 1);
  bool head = (lane == 0) || (prev != key);
  uint32_t heads = __ballot_sync(0xffffffffu, head);
  uint32_t run_start = 31u - __clz(heads & (0xffffffffu >> (31u - lane)));
  uint32_t above = (lane == 31u) ? 0u : (heads >> (lane + 1u));
  uint32_t run_last = above ? (lane + (uint32_t)__ffs(above) - 1u) : 31u;
  uint32_t width = 1u << steps, pos = (lane - run_start) & (width - 1u);
  uint32_t chunk_last = min(run_last, lane - pos + width - 1u);
  float ax = 1.f - c.wx, ay = 1.f - c.wy, az = 1.f - c.wz;
  float v[16];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float w = ((i & 4) ? c.wx : ax) * (((i & 2) ? c.wy : ay) * ((i & 1) ? c.wz : az));
    v[2 * i] = live ? g.x * w : 0.f;
    v[2 * i + 1] = live ? g.y * w : 0.f;
  }
  for (uint32_t d = 1; d < width; d <<= 1) {
    bool take = lane + d <= chunk_last;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      float o = __shfl_down_sync(0xffffffffu, v[i], d);
      v[i] += take ? o : 0.f;
    }
  }
  if (!live || pos != 0u) return;
  uint32_t rows[8];
  corner_rows(c, mask, base, rows);
  if ((c.ix & 1u) == 0u) {
    float4* t4 = reinterpret_cast<float4*>(gtable);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bool odd = rows[i] & 1u;
      float4 q = odd ? make_float4(v[2 * i + 8], v[2 * i + 9], v[2 * i], v[2 * i + 1]) : make_float4(v[2 * i], v[2 * i + 1], v[2 * i + 8], v[2 * i + 9]);
      red_add(t4 + (rows[i] >> 1), q);
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) red_add(gtable + rows[i], make_float2(v[2 * i], v[2 * i + 1]));
}

}  // namespace acn
