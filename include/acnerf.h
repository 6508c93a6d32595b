/* acnerf.h - C ABI of libacnerf_sm100a.so: the B200 (sm_100a) per-ray hot path of
 * Adaptive City NeRF (psklavos1/NeRF-Sys).
 *
 * Every entry point replaces one piece of the reference's PyTorch path; the
 * reference interface is cited as file:line under /root/reference/adaptive_nerf/.
 * The reference has no FFI of its own (it is pure Python): this header is what a
 * ctypes / torch custom-op binding on the reference side binds.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name ends in _host;
 *  - the caller owns all memory, including workspaces (sizes via acn_*_ws_bytes);
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*),
 *    re-entrant, and holds no global device state;
 *  - return value: 0 = ok, 1 = invalid argument, 2 = CUDA error, 3 = unsupported
 *    device; acn_last_error() returns a thread-local message. Nothing throws
 *    across the ABI, nothing calls exit();
 *  - there is no CPU path: a NULL/host pointer where a device pointer is
 *    expected is an error, not a fallback.
 *  - all tensors fp32 row-major unless stated; indices int32; masks uint8.
 */
#ifndef ACNERF_H_
#define ACNERF_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built -fvisibility=hidden; only this header is exported */
#endif

#define ACN_OK 0
#define ACN_EINVAL 1
#define ACN_ECUDA 2
#define ACN_EDEVICE 3

#define ACN_MAX_LEVELS 32
#define ACN_MAX_EXPERTS 16 /* experts per launch (kernel-parameter table) */
#define ACN_TILE 128       /* pair segments are padded to this many slots */

/* ---- library ------------------------------------------------------------ */
int acn_version(void);                  /* ABI version, currently 1 */
const char* acn_last_error(void);       /* thread-local, never NULL */
int acn_device_ok(int device);          /* ACN_OK iff `device` is compute capability 10.x */
int acn_num_sms(int device);            /* multiprocessor count (148 on B200), <0 on error */
/* Launch-sizing hints of the CALLING THREAD (no device state; launches read them on the host at launch time).  Returns
 * the previous value, -1 for an unknown key.  Tuning knobs only: results are identical for every value.
 *   ACN_HINT_GATHER_LEVELS = 1, 2, 4, 8 or 16: levels of the hash grid one thread of acn_ngp_encode_fwd walks
 *   (0 = default: 8 for lists of >= 4 Mi slots, else 4). */
#define ACN_HINT_GATHER_LEVELS 1
#define ACN_HINT_COUNT 4
int acn_launch_hint(int key, int value);

/* ---- descriptors -------------------------------------------------------- */
/* models/encodings.py:180-286 (HashGridEncoder ctor): levels, features_per_level,
 * log2_hashmap_size, interpolation, and the per-level resolutions the reference
 * computes as floor(min_res * g^l) in fp32 (encodings.py:203-215); the host side
 * computes them with the same torch expression and passes them here. */
typedef struct AcnGridDesc {
  int32_t levels;   /* L <= 32 */
  int32_t features; /* F, must be 2 */
  int32_t log2_T;   /* rows per level = 1 << log2_T */
  int32_t interp;   /* 0 Linear, 1 Smoothstep, 2 Nearest */
  int32_t res[ACN_MAX_LEVELS];
} AcnGridDesc;

/* Where sample points come from.  Either explicit (P,6) [xyz,dir] rows
 * (MetaNGP.forward's argument, meta_ngp.py:226-241) or implicit: point
 * p = ray (p / S), sample (p % S), xyz = o + d*t, dir = d (ray_rendering.py:317-319),
 * which never materialises the (N*S,6) tensor. */
typedef struct AcnPoints {
  const float* x6;   /* (P,6) or NULL */
  const float* rays; /* (N,8) [o,d,near,far] when x6 == NULL */
  const float* t;    /* (N,S) */
  int32_t S;
  int64_t P;         /* number of points */
  const float* t_T;  /* optional (S,N) copy of t, N = P/S.  Non-NULL selects SAMPLE-MAJOR point ids: point p is sample
                        p / N of ray p % N and its t is t_T[p]; everything indexed by point id (pair lists, outputs) is
                        then in (S,N) order.  A renderer's choice for rays that are neighbouring pixels: consecutive
                        points - the lanes of every kernel - are the same sample of adjacent pixels and share grid
                        cells, where consecutive samples of one ray share nothing above res ~32. */
} AcnPoints;

/* Routed (point, expert) pairs, grouped by expert, each group padded to a
 * multiple of ACN_TILE slots.  All-NULL = identity: one expert, slot m == point m,
 * weight 1 (the active_module path, meta_container.py:302-304). */
typedef struct AcnPairs {
  const int32_t* point; /* (Mcap,) point id, -1 = padding */
  const float* weight;  /* (Mcap,) routing weight; NULL with point != NULL = every weight is 1 and no point id occurs twice
                           (an owner's received rows under expert parallelism): results are stored, not accumulated */
  const int32_t* seg;   /* (E+1,) segment starts (multiples of ACN_TILE); seg[E] = used slots */
  int64_t Mcap;         /* capacity of the pair arrays (identity: == P) */
} AcnPairs;

/* One hash-grid expert (MetaNGP, meta_ngp.py:21-142) in the production shape
 * L*F=32 -> 64 -> 64 -> {sigma 1, geo 15}; [geo15|SH16]=31 -> 64 -> 64 -> 3.
 * Weights are nn.Linear layout (out,in), i.e. exactly the reference's
 * state_dict tensors or the caller's fast-weight overrides (metamodule.py:129-156).
 * layer order: 0 sigma_trunk.0, 1 sigma_trunk.1, 2 sigma_head, 3 geo_head,
 *              4 color_mlp.0, 5 color_mlp.1, 6 color_mlp.2 */
#define ACN_NGP_LAYERS 7
typedef struct AcnNgpExpert {
  const float* table;              /* (L*T, 2) */
  const float* w[ACN_NGP_LAYERS];
  const float* b[ACN_NGP_LAYERS];
  float box_min[3];                /* expert's own AABB (meta_ngp.py:155-158) */
  float extent[3];
} AcnNgpExpert;

typedef struct AcnNgpExpertGrads { /* outputs of the backward pass; NULL = not wanted */
  float* table;                    /* (L*T, 2), ACCUMULATED into (caller zeroes) */
  float* w[ACN_NGP_LAYERS];        /* overwritten */
  float* b[ACN_NGP_LAYERS];
} AcnNgpExpertGrads;

/* ---- a1-a3: rays -------------------------------------------------------- */
/* nerfs/ray_sampling.py:111-136 get_ray_directions -> dirs (H,W,3) */
int acn_ray_directions(int H, int W, float fx, float fy, float cx, float cy, int center_pixels,
                       float* dirs, void* stream);
/* nerfs/ray_sampling.py:50-108 get_rays (+ scene_box.py:45-107 ray_aabb_intersect):
 * dirs (n,3) camera frame, c2w_host 12 floats row-major (3,4); aabb_host 6 floats
 * [min,max] or NULL (then near/far constants) -> rays (n,8). */
int acn_get_rays(const float* dirs, int64_t n, const float* c2w_host, const float* aabb_host,
                 float near_c, float far_c, float aabb_max_bound, float aabb_invalid_value,
                 float* rays, void* stream);
/* fused a1+a2 for a full frame; rows [row0,row0+rows) of an HxW image (ray-sharded render) */
int acn_raygen(int H, int W, int row0, int rows, float fx, float fy, float cx, float cy, int center_pixels,
               const float* c2w_host, const float* aabb_host, float near_c, float far_c,
               float* rays, void* stream);
/* scene_box.py:45-107 on explicit origins/directions */
int acn_ray_aabb(const float* o, const float* d, int64_t n, const float* aabb_host, float max_bound,
                 float invalid_value, float* near_out, float* far_out, void* stream);
/* nerfs/ray_sampling.py:139-176 clamp_rays_near_far.  has_near/has_far select the
 * override bounds; override_none!=0 = the `near_far_override=None` branch (rays untouched). */
int acn_clamp_near_far(const float* rays_in, int64_t n, int override_none, int has_near, float near_v,
                       int has_far, float far_v, float eps, float invalid_value, float* rays_out,
                       uint8_t* valid, void* stream);

/* ---- a4-a5: stratified sampling ----------------------------------------- */
/* nerfs/ray_rendering.py:262-287.  tau = torch.linspace(0,1,S) values (S,), u (N,S)
 * jitter in [0,1) or NULL (eval).  t (N,S). */
int acn_stratified_t(const float* rays, int64_t N, int S, const float* tau, const float* u, float* t,
                     void* stream);
/* Proposal (hierarchical) resampling - ADDITIVE: the reference has no proposal sampler (SURVEY headline 4; north_star names
 * one).  t, weights (N,S) from a coarse pass (3 <= S <= 256), u (N,S2) in [0,1) ASCENDING per ray.  t_new (N,S2): inverse-CDF
 * samples of the piecewise-constant pdf over the midpoint bins (weights[1..S-2] + 1e-5); t_merged (N,S+S2) or NULL: the
 * coarse and the new distances merged in ascending order. */
int acn_sample_pdf(const float* t, const float* weights, const float* u, int64_t N, int S, int S2, float* t_new,
                   float* t_merged, void* stream);
/* nerfs/ray_rendering.py:317-319: (N,8),(N,S) -> x6 (N*S,6).  API parity / tests only;
 * the render path never materialises this. */
int acn_sample_points(const float* rays, const float* t, int64_t N, int S, float* x6, void* stream);

/* ---- a6-a7: routing ------------------------------------------------------ */
/* models/inr/meta_container.py:97-134 _routing.  centroids_host (K,3) HOST floats (they
 * travel in the kernel-parameter block).  margin>1: weights (P,K) written, hard
 * untouched; else hard (P,) int64 argmin written. */
int acn_routing(const AcnPoints* pts, const float* centroids_host, int K, float margin, int cluster_2d,
                float* weights, int64_t* hard, void* stream);
/* Fused routing + stable pack-by-expert (replaces K x nonzero/index_select,
 * meta_container.py:305-337).  Outputs the pair arrays (capacity Mcap slots; a
 * speculative size comes from acn_route_pack_capacity) and seg (K+1).  Nothing is
 * written past Mcap; seg[K] always reports the slots NEEDED, so a caller that reads
 * seg[K] > Mcap re-runs with larger arrays.  ws: acn_route_pack_ws_bytes(P,K) bytes of
 * device memory, 16-byte aligned (per-block counts, one decision word per point and a
 * weight row per point for the samples with more than one member: 4 + 4*KT bytes per
 * point, KT = K rounded up to 2/4/8/16). */
int64_t acn_route_pack_capacity(int64_t P, int K);
int64_t acn_route_pack_ws_bytes(int64_t P, int K);
int acn_route_pack(const AcnPoints* pts, const float* centroids_host, int K, float margin, int cluster_2d,
                   int32_t* pair_point, float* pair_weight, int32_t* seg, int64_t Mcap, void* ws,
                   void* stream);

/* Device-side reading of a segment table for steps that never read it back on the host: live[k] = (seg[k+1] > seg[k])
 * (expert k received a sample), *status |= 1 when seg[K] > Mcap (the pair list was truncated at its speculative capacity;
 * sticky - the host polls it every so often). */
int acn_route_status(const int32_t* seg, int K, int64_t Mcap, int32_t* live, int32_t* status, void* stream);

/* ---- a8-a11: encoders ---------------------------------------------------- */
/* models/encodings.py:331-358: table row ids of the 8 corners, idx (P,L,8) int32,
 * corner c = 4*bx+2*by+bz, level offset included (bit-exact contract). */
int acn_hashgrid_indices(const float* x01, int64_t P, const AcnGridDesc* g, int32_t* idx, void* stream);
/* HashGridEncoder.forward (encodings.py:293-305, 331-381): x01 (P,3) -> y (P,L*F) */
int acn_hashgrid_fwd(const float* x01, int64_t P, const float* table, const AcnGridDesc* g, float* y,
                     void* stream);
/* its autograd backward: grad_table (L*T,F) += scatter(dy); caller zeroes grad_table */
int acn_hashgrid_bwd(const float* x01, int64_t P, const float* dy, const AcnGridDesc* g,
                     float* grad_table, void* stream);
/* SHEncoder.forward (encodings.py:27-81,133-151): d (P,3) -> (P,16) */
int acn_sh16(const float* d, int64_t P, float* out, void* stream);
/* FrequencyEncoder.torch_forward (encodings.py:437-444): x (P,3) -> (P, 3 + 6*n_freq) or (P,6*n_freq) */
int acn_freq_encode(const float* x, int64_t P, int n_freq, int include_input, float* out, void* stream);

/* ---- a8-a14 fused: hash-grid experts over routed pairs ------------------- */
/* The expert evaluation is two stream-ordered stages that exchange the encoding in a
 * feature-major buffer enc (32, ld), ld = acn_ngp_ld(Mcap) (slot index contiguous):
 *   encode: world -> unit cube of the expert's own box -> hash grid   (meta_ngp.py:155-163)
 *   mlp   : sigma trunk, heads, trunc_exp, SH16, colour MLP, sigmoid  (meta_ngp.py:171-241)
 *           and the container's combine out[p] (+)= w*y                (meta_container.py:305-321)
 * Two stages so that autograd can ask for the MLP gradients alone (the meta inner loop,
 * meta_core.py:54-59) without paying for the hash-table scatter. */
int64_t acn_ngp_ld(int64_t Mcap);
int64_t acn_ngp_save_bytes(int64_t Mcap);
int64_t acn_ngp_bwd_ws_bytes(int64_t Mcap, int E);
int acn_ngp_encode_fwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                       const AcnGridDesc* g, float* enc, void* stream);
/* grads[e].table (L*T,2) is ACCUMULATED into (caller zeroes); NULL = skip that expert */
int acn_ngp_encode_bwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                       const AcnGridDesc* g, const float* d_enc, const AcnNgpExpertGrads* grads, void* stream);
/* out (P,4) [rgb,sigma]: identity pairs -> plain store; routed -> out[p] += w*y (caller
 * zeroes out).  save: acn_ngp_save_bytes(Mcap) bytes kept for backward, or NULL (inference). */
int acn_ngp_mlp_fwd(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                    const float* enc, float* out, void* save, void* stream);
/* d_out (P,4) -> grads[e].w/b (overwritten; NULL = not wanted) and d_enc (32, ld).
 * ws: acn_ngp_bwd_ws_bytes(Mcap,E). */
int acn_ngp_mlp_bwd(const AcnPairs* pairs, const AcnNgpExpert* experts, int E, const float* enc,
                    const float* d_out, const void* save, const AcnNgpExpertGrads* grads, float* d_enc,
                    void* ws, void* stream);

/* ---- a11-a14 on the tensor cores (tcgen05 / TMEM) --------------------------- */
/* Same contract as acn_ngp_mlp_fwd / _bwd (meta_ngp.py:171-241, meta_container.py:305-321), computed
 * with fp16 operands and fp32 accumulation on the 5th-generation tensor cores - the precision class of
 * the reference's own GPU path (TF32 matmuls, nerf_runner.py:42-43; fp16 autocast, meta_core.py:38).
 * The backward recomputes the forward per 128-slot tile, so no activations are saved: it takes the
 * points (for the view directions) and the encoding again.  ws: acn_ngp_bwd_tc_ws_bytes().
 * grads[e].w/b are overwritten (NULL = not wanted); d_enc (32, ld) as in acn_ngp_mlp_bwd. */
int acn_ngp_mlp_fwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                       const float* enc, float* out, void* stream);
/* The same forward with the hash-grid gather inside the kernel (models/inr/meta_ngp.py:192-241 -> models/encodings.py:331-381
 * in one launch): experts[e].table is read directly, the (32, ld) encoding is never written.  Inference only (no
 * activations are kept); interpolation "linear" or "smoothstep".  out as acn_ngp_mlp_fwd_tc. */
int acn_ngp_fused_fwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                         const AcnGridDesc* g, float* out, void* stream);
int64_t acn_ngp_bwd_tc_ws_bytes(void);
int acn_ngp_mlp_bwd_tc(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                       const float* enc, const float* d_out, const AcnNgpExpertGrads* grads, float* d_enc,
                       void* ws, void* stream);
/* Training backward of the whole expert stage (MLP backward + hash-table scatter) as a software pipeline over
 * n_chunks (1..8) chunks of the pair list: chunk c's scatter (acn_ngp_encode_bwd semantics, grads[e].table
 * accumulated, caller zeroes) runs on side_stream while chunk c+1's MLP backward runs on stream, on the same SMs;
 * `stream` is joined with the side stream before return.  ws: acn_ngp_bwd_tc_pipelined_ws_bytes(n_chunks). */
int64_t acn_ngp_bwd_tc_pipelined_ws_bytes(int n_chunks);
int acn_ngp_bwd_tc_pipelined(const AcnPoints* pts, const AcnPairs* pairs, const AcnNgpExpert* experts, int E,
                             const AcnGridDesc* g, const float* enc, const float* d_out,
                             const AcnNgpExpertGrads* grads, float* d_enc, void* ws, int n_chunks, void* stream,
                             void* side_stream);
/* Descriptor self-test of the three shared-memory operand layouts the kernels above use
 * (X (128,64), Y (64,64), G (128,64) fp32 in; D1 = X Y^T (128,64), D2 = G^T X (64,64), D3 = G Y (128,64));
 * G is rounded to bf16 when mixed != 0, everything else to fp16. */
int acn_tc_selftest(const float* X, const float* Y, const float* G, int mixed, float* D1, float* D2,
                    float* D3, void* stream);

/* ---- a15: positional-encoding MLP expert (MetaNeRF) as tcgen05 GEMM layers ----------------- */
/* models/inr/meta_vanilla.py:109-154 (config 1).  Activations are fp16 "chunk-major" tensors in device memory:
 * (chunks, ld, 8) halfs, feature f of row m at ((f/8)*ld + m)*8 + f%8, ld a multiple of 128 >= rows.
 * The host composes the network (any width / depth / skips) from these calls; weights are the reference's
 * nn.Linear tensors (out,in) fp32, converted to fp16 operands inside the kernels, fp32 accumulation. */
/* [prefix (n_prefix fp32 columns of `prefix`, or none) | FrequencyEncoder(xyz if which==0 else dir)]
 * (encodings.py:437-444) -> chunks [chunk0, chunk0 + ceil(width/8)) of out_h, zero padded. */
int acn_pe_encode_h(const AcnPoints* pts, int which, int n_freq, int include_input, const float* prefix,
                    int64_t prefix_ld, int n_prefix, void* out_h, int64_t ld, int chunk0, void* stream);
/* Y = act(A W^T + b) [zeroed where mask feature <= 0]: A = k_chunks chunks of a_h (k_valid real features),
 * W element (n,k) at W[n*ldw+k] (transposed != 0: W[k*ldw+n], the dgrad form), act 0 identity / 1 ReLU.
 * Outputs: chunk-major fp16 (feature n -> chunk out_chunk0 + n/8 if < out_chunks) and/or fp32 rows (M, out_f_ld). */
int acn_dense_tc(const void* a_h, int k_chunks, int64_t M, int64_t ld, const float* W, int64_t ldw, int transposed,
                 const float* bias, int n_out, int k_valid, int act, const void* mask_h, int64_t mask_ld,
                 int mask_chunks, void* out_h, int64_t out_ld, int out_chunk0, int out_chunks, float* out_f,
                 int64_t out_f_ld, void* stream);
/* dW (n_out,k_valid) = dZ^T A (row stride ldw), db = column sums of dZ, both times *inv_scale_dev if given;
 * accumulated in TMEM per CTA and reduced in a fixed order.  ws: acn_wgrad_tc_ws_bytes(n_out, k_chunks). */
int64_t acn_wgrad_tc_ws_bytes(int n_out, int k_chunks);
int acn_wgrad_tc(const void* dz_h, int dz_chunks, const void* a_h, int k_chunks, int64_t M, int64_t ld, int n_out,
                 int k_valid, float* dW, int64_t ldw, float* db, const float* inv_scale_dev, void* ws, void* stream);
/* out (P,4) = [sigmoid?(rgb_raw[:, :3]), trunc_exp(heads[:, sigma_col])] */
int acn_pe_final_fwd(const float* rgb_raw, int64_t rgb_ld, const float* heads, int64_t heads_ld, int sigma_col,
                     int use_sigmoid, int64_t P, float* out, void* stream);
/* d_out (P,4), out (P,4) -> g_rgb_h (2 chunks: the scaled gradient of the 3 colour logits), dsraw (P,) scaled
 * d sigma_raw, scale4 = [s, 1/s, max|d_out|, -] with s the power of two putting max|d_out| near 2^10. */
int acn_pe_final_bwd(const float* d_out, const float* out, int use_sigmoid, int64_t P, void* g_rgb_h, int64_t ld,
                     float* dsraw, float* scale4, void* stream);
/* feature `feature` of a chunk-major fp16 tensor <- fp32 vector v (P,) */
int acn_set_feature_h(void* t_h, int64_t ld, int feature, const float* v, int64_t P, void* stream);

/* ---- a16: shared background head ----------------------------------------- */
/* meta_container.py:347-382: d (N,3) taken at `stride` floats (stride 8 reads rays[:,3:6]
 * in place) -> rgb (N,3).  w0 (32,16) b0 (32) w1 (3,32) b1 (3). */
int64_t acn_bg_save_bytes(int64_t N);
int acn_bg_forward(const float* d, int64_t stride, int64_t N, const float* w0, const float* b0,
                   const float* w1, const float* b1, float* rgb, void* save, void* stream);
int64_t acn_bg_bwd_ws_bytes(int64_t N);
int acn_bg_backward(const float* d_rgb, int64_t N, const float* w0, const float* w1, const void* save,
                    float* g_w0, float* g_b0, float* g_w1, float* g_b1, void* ws, void* stream);

/* ---- a17: alpha compositing ---------------------------------------------- */
/* nerfs/ray_rendering.py:114-165 volume_render.  rgb_sigma (N,S,4), t (N,S), bg (N,3) or NULL
 * -> rgb (N,3), depth (N), weights (N,S) or NULL, acc (N).  No early termination. */
int acn_composite_fwd(const float* rgb_sigma, const float* t, const float* bg, int64_t N, int S,
                      float sigma_scale, float* rgb, float* depth, float* weights, float* acc, void* stream);
/* acn_composite_fwd with the (rgb,sigma) row of (ray r, sample s) at rgb_sigma + 4*(r*ray_stride + s*sample_stride):
 * (S, 1) is the reference's (N,S,4) layout, (1, N) the (S,N,4) output of a sample-major point set (AcnPoints.t_T).
 * t, weights stay (N,S). */
int acn_composite_fwd_strided(const float* rgb_sigma, int64_t ray_stride, int64_t sample_stride, const float* t,
                              const float* bg, int64_t N, int S, float sigma_scale, float* rgb, float* depth,
                              float* weights, float* acc, void* stream);
/* The same for a sample-major point set: rgb_sigma_T (S,N,4), t_T (S,N) -> rgb (N,3), depth (N), weights_T (S,N) or NULL, acc (N);
 * one thread per ray, sequential transmittance product (the order torch.cumprod uses), fully coalesced rows. */
int acn_composite_fwd_sm(const float* rgb_sigma_T, const float* t_T, const float* bg, int64_t N, int S, float sigma_scale,
                         float* rgb, float* depth, float* weights_T, float* acc, void* stream);
/* its backward: g_rgb (N,3), g_depth (N) or NULL, g_acc (N) or NULL, g_weights (N,S) or NULL
 * -> d_rgb_sigma (N,S,4), d_bg (N,3) or NULL */
int acn_composite_bwd(const float* rgb_sigma, const float* t, const float* bg, int64_t N, int S,
                      float sigma_scale, const float* g_rgb, const float* g_depth, const float* g_acc,
                      const float* g_weights, float* d_rgb_sigma, float* d_bg, void* stream);

/* ---- a18: photometric loss ------------------------------------------------ */
/* nerfs/losses.py:10-32 + color_space.py:22-66.  space: 0 linear, 1 srgb, 2 identity.
 * loss (1,) = mean((pred'-gt')^2); d_pred (N,3) = dloss/dpred (unit upstream grad) or NULL.
 * ws: 2048 floats. */
int acn_mse_loss(const float* pred, const float* gt, int64_t N, int space, float* loss, float* d_pred,
                 float* ws, void* stream);

/* The same loss with NO caller workspace and NO allocation of any kind (no cudaMalloc / cudaFree, no allocating library
 * routine), for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
 * No counterpart in the reference.  The per-CTA partial sums live in a
 * file-scope `static __device__` array of ACN_MSE_STATIC_PARTIALS floats (8 KiB, held by every GPU that loads the library
 * until the process exits); N above ACN_MSE_STATIC_MAX_RAYS is rejected with ACN_EINVAL.  Same kernels, grid and summation
 * order as acn_mse_loss: results are bit-identical to it.  The array is shared by every call and keeps the last call's
 * contents (each call overwrites all the entries it reads, nothing needs clearing); two calls that OVERLAP on the GPU
 * (different streams, or different threads on one device) race on it - order them on one stream or with events. */
#define ACN_MSE_STATIC_PARTIALS 2048
#define ACN_MSE_STATIC_MAX_RAYS (1 << 24)
int acn_mse_loss_static(const float* pred, const float* gt, int64_t N, int space, float* loss, float* d_pred,
                        void* stream);

/* ---- (f)2: fused optimiser ------------------------------------------------ */
/* torch.optim.Adam single-tensor math (common/utils.py:16-76 builds torch Adam):
 * step is the 1-based step count; grad_scale multiplies the gradient first
 * (clip coefficient and/or 1/loss-scale). */
int acn_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                  float beta1, float beta2, float eps, float weight_decay, int64_t step,
                  const float* grad_scale_dev, void* stream);
/* fast-weight SGD step of the inner loop (pipelines/offline_stage/meta_core.py:61-64): out = w - lr*g */
int acn_sgd_axpy(const float* w, const float* g, float lr, int64_t n, float* out, void* stream);
/* sum of squares of a tensor accumulated into acc (1,) (clip_grad_norm_ building block) */
int acn_sumsq_accum(const float* x, int64_t n, float* acc, void* stream);

/* Multi-tensor forms: the whole parameter set in one launch each (a 15-tensor expert plus its 128 MiB table
 * otherwise costs ~40 launches per step).  Step counts live on the device (steps_dev[step_index], int32,
 * incremented by the call) so the step can be replayed from a CUDA graph; bias corrections are evaluated in
 * double like torch's.  The gradient multiplier (clip_grad_norm_ + GradScaler.unscale_ semantics, meta_core.py:123-141,
 * 181-190) is written to coef_dev when write_coef (first chunk of a step) and applied to every gradient. */
#define ACN_ADAM_MAX_TENSORS 64
typedef struct AcnAdamTensor {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t n;
  float lr;
  float weight_decay;
  int32_t step_index;
  int32_t flags;   /* low 8 bits: liveness slot + 1 (0 = always live); ACN_ADAM_ZERO_GRAD: zero the gradient once it is read */
} AcnAdamTensor;
#define ACN_ADAM_ZERO_GRAD 0x100
/* live_dev (int32, may be NULL): a tensor with liveness slot s is processed only when live_dev[s] != 0 - the device-side
 * form of "p.grad is None -> torch.optim.Adam skips the parameter" (an expert that received no sample this step,
 * meta_container.py:309-311), so that a step whose routing is never read back by the host can live in a CUDA graph.
 * found_inf_dev (float, may be NULL): GradScaler's flag; != 0 skips the whole step (runtime_adapt.py:267).
 * coef_dev (2 floats): [0] = gradient multiplier = grad_scale * min(1, max_norm / (sqrt(sumsq) * |grad_scale| + 1e-6)),
 * i.e. the clip acts on the UNSCALED norm as scaler.unscale_ + clip_grad_norm_ do; [1] != 0 = step skipped (found_inf,
 * or a non-finite norm). */
int acn_sumsq_multi(const AcnAdamTensor* tensors_host, int count, const int32_t* live_dev, float* acc, void* stream);
int acn_adam_multi(const AcnAdamTensor* tensors_host, int count, float beta1, float beta2, float eps,
                   int32_t* steps_dev, const float* sumsq_dev, float max_norm, const float* grad_scale_dev,
                   const float* found_inf_dev, const int32_t* live_dev, float* coef_dev, int write_coef, void* stream);

/* ---- (e): expert-parallel dispatch / combine around the sample all-to-all -
 * New design (the reference is single-GPU); replaces, across GPUs, the per-expert
 * nonzero / index_select / index_add_ loop of models/inr/meta_container.py:305-337.
 * acn_a2a_pack: compacts a routed pair list (drops the -1 padding) into send rows
 *   rows6 (M,6) = [xyz, dir] built from the implicit point form, grouped by expert
 *   (= grouped by destination rank under block placement); out_point/out_weight (M,)
 *   are the point id and routing weight of each row, counts (K+1,) the rows per expert
 *   (counts[K] = M), offs (K+1,) their exclusive prefix.  Buffers are sized for
 *   pairs->Mcap rows by the caller.
 * acn_a2a_combine: out[point[j]] += weight[j] * y[j] for returned [rgb,sigma] rows
 *   (out (P,4), caller zeroes);  acn_a2a_combine_bwd: d_y[j] = weight[j] * d_out[point[j]]. */
int acn_a2a_pack(const AcnPoints* pts, const AcnPairs* pairs, int K, float* rows6, int32_t* out_point,
                 float* out_weight, int32_t* counts, int32_t* offs, void* stream);
int acn_a2a_combine(const float* y, const int32_t* point, const float* weight, int64_t M, float* out, void* stream);
int acn_a2a_combine_bwd(const float* d_out, const int32_t* point, const float* weight, int64_t M, float* d_y,
                        void* stream);

/* The same exchange with the rows living in the OWNER's memory, written / read by the kernels themselves over
 * NVLink peer mappings instead of travelling through a collective.  AcnA2aRoute: base[e] = the buffer of expert e's
 * owner as mapped into this process (a local pointer when this rank owns e), row0[e] = first row of this rank's block
 * of expert e in it; pair j of the compact list (expert e = max{k: offs[k] <= j}) is row row0[e] + j - offs[e].
 *   acn_a2a_count        counts (K+1,), offs (K+1,) of a routed pair list (first half of acn_a2a_pack); counts[K] = -1
 *                        when seg[K] > pairs->Mcap, i.e. the list was truncated at its speculative capacity
 *   acn_a2a_push_rows    writes the [xyz,dir] rows (24 B) into the owners' row buffers; out_point/out_weight as acn_a2a_pack
 *   acn_a2a_pull_combine out[point[j]] += weight[j] * y_owner[row]   (16-byte loads from the owners' result buffers)
 *   acn_a2a_push_grad    d_y_owner[row] = weight[j] * d_out[point[j]] (16-byte stores into the owners' gradient buffers)
 * Ordering between ranks (rows landed before the owner evaluates, results complete before they are pulled, buffers free
 * before they are rewritten) is the caller's: a stream-ordered barrier between the phases. */
typedef struct AcnA2aRoute {
  void* base[ACN_MAX_EXPERTS];
  int64_t row0[ACN_MAX_EXPERTS];
} AcnA2aRoute;
int acn_a2a_count(const AcnPairs* pairs, int K, int32_t* counts, int32_t* offs, void* stream);
/* row0_dev (K int64, may be NULL) replaces route->row0 with values a kernel computed (acn_p2p_plan): no host read between
 * routing and dispatch.  cap_rows > 0: rows at or beyond it are not written (an overflowing owner slab is reported by
 * acn_p2p_plan's status word instead of being overrun).  M in pull_combine / push_grad is an upper bound; the live pair
 * count is offs[K] on the device. */
int acn_a2a_push_rows(const AcnPoints* pts, const AcnPairs* pairs, int K, const int32_t* offs, const AcnA2aRoute* route,
                      const int64_t* row0_dev, int64_t cap_rows, int32_t* out_point, float* out_weight, void* stream);
int acn_a2a_pull_combine(const AcnA2aRoute* route, const int64_t* row0_dev, const int32_t* point, const float* weight,
                         const int32_t* offs, int K, int64_t M, float* out, void* stream);
int acn_a2a_push_grad(const float* d_out, const int32_t* point, const float* weight, const int32_t* offs, int K, int64_t M,
                      const AcnA2aRoute* route, const int64_t* row0_dev, int64_t cap_rows, void* stream);
/* Exchange slabs (one process per GPU, CUDA IPC): acn_p2p_alloc = cudaMalloc + a 64-byte handle the peers pass to
 * acn_p2p_open (maps the slab with peer access); the host layer owns both and releases them with _close / _free. */
#define ACN_MAX_PEERS 16
#define ACN_P2P_PAYLOAD_FLOATS 1024
/* head of every slab: int32 flag[16] | status | epoch | pad to 32 ints | table[16][ACN_MAX_EXPERTS+1] | pad to 4096 B |
 * float payload[2][16][ACN_P2P_PAYLOAD_FLOATS] (two buffers alternating with the barrier epoch) */
#define ACN_P2P_SYNC_BYTES (4096 + 2 * ACN_MAX_PEERS * ACN_P2P_PAYLOAD_FLOATS * 4)
int acn_p2p_alloc(int64_t bytes, void** ptr, void* handle64); /* the whole slab is zeroed */
int acn_p2p_free(void* ptr);
int acn_p2p_open(const void* handle64, void** ptr);
int acn_p2p_close(void* ptr);
/* The next barrier between the `world` ranks whose slabs are sync_bases_host[0..world-1] (as mapped into this process),
 * stream-ordered: the peer stores of this rank's earlier kernels are visible to a rank once that rank's barrier kernel
 * has returned.  The barrier number lives in the rank's own slab and is advanced by the kernel, so the launch carries no
 * per-call state and can be replayed from a CUDA graph; every rank must issue the same sequence of barriers.
 *  counts  != NULL: the rank first writes its n_counts ints into row `rank` of every slab's table (an all-gather);
 *  payload != NULL: the rank writes n_payload (<= ACN_P2P_PAYLOAD_FLOATS) floats into its slot of every slab, and after
 *                   the wait every rank sums the world slots in rank order into reduced (n_payload floats) times scale:
 *                   a deterministic small all-reduce riding on the barrier (background-head gradients, clip norms).
 * A peer that does not arrive within timeout_s seconds sets bit 0 of the status word (int 16 of the sync area) and
 * traps the kernel, so that nothing downstream consumes a half-written slab; the CUDA error surfaces at the host's next
 * synchronisation. */
int acn_p2p_barrier(void* const* sync_bases_host, int world, int rank, const int32_t* counts, int n_counts,
                    const float* payload, int n_payload, float* reduced, float scale, float timeout_s, void* stream);
/* The barrier in two launches: ARRIVE publishes the payload and raises this rank's flag without waiting; WAIT blocks until every
 * rank's flag of that epoch is up, then reduces the payload slots.  Kernels enqueued between the two absorb the skew between the
 * ranks (gradients that are complete early in a backward pass).  Each arrive must be followed by its wait before any other
 * barrier on the same slabs. */
int acn_p2p_arrive(void* const* sync_bases_host, int world, int rank, const float* payload, int n_payload, void* stream);
int acn_p2p_wait(void* const* sync_bases_host, int world, int rank, int n_payload, float* reduced, float scale, float timeout_s,
                 void* stream);
/* Device-side layout of one exchange (parallel.peer_layout on the GPU).  Reads the count table of this rank's slab
 * (filled by the counts barrier); owner_host[e] = rank owning expert e.  Layout of an owner's slab: its experts in
 * order, each expert's block = the sub-blocks of source ranks 0..world-1 back to back, blocks padded to ACN_TILE rows.
 *   row0_dev (K int64)       first row of THIS rank's sub-block of expert e in its owner's slab
 *   seg_own (n_own+1 int32)  segment table of this rank's own experts over its slab rows
 *   point_own (cap_rows)     row id for live rows, -1 for the padding rows of each block (an AcnPairs point list)
 *   live_own (n_own int32)   expert received a sample from any rank
 *   g_rows                   this rank's gradient rows (cap_rows,4): the padding rows are zeroed
 *   status |= 2 when some owner needs more than cap_rows rows, |= 4 when a rank's pair list was truncated (sticky). */
int acn_p2p_plan(const void* sync_base_own, int world, int rank, int K, const int32_t* owner_host, int64_t cap_rows,
                 int64_t* row0_dev, int32_t* seg_own, int32_t* point_own, int32_t* live_own, float* g_rows,
                 int32_t* status, void* stream);

/* ---- a21 / (f)-1: occupancy-guided rendering (nerfs/ray_rendering.py:349-558; models/inr/meta_ngp.py:389-443) ----------
 * The reference calls nerfacc 0.5.3 (OccGridEstimator.sampling, render_weight_from_density, accumulate_along_rays) and merges
 * the experts' segments in a Python loop over rays; these entry points are what replaces both.  nerfacc is not vendored and
 * absent here: the traversal follows its documented semantics, restated in oracle/acnerf_oracle.py (parity unpinned).
 * Grid: binaries (levels, R, R, R) uint8, cell = (ix*R + iy)*R + iz; aabbs_host (levels, 6) host floats, level l = the region
 * of interest scaled by 2^l about its centre (OccGridEstimator.aabbs).  Packed lists: offsets (N,) int64 = first segment of
 * ray r, counts (N,) int32. */
#define ACN_OCC_MAX_LEVELS 8
/* pass 1: counts[r] = samples of ray r (<= max_samples).  near_jitter (N,) in [0,1) or NULL: near += jitter * step (training) */
int acn_occ_march_count(const float* rays, const float* near_jitter, int64_t N, const uint8_t* binaries, const float* aabbs_host,
                        int levels, int R, float step, float cone_angle, int max_samples, int32_t* counts, void* stream);
/* pass 2: offsets = exclusive scan of counts; writes ray_idx / t0 / t1 (and, if not NULL, cells = level*R^3 + cell per sample) */
int acn_occ_march_write(const float* rays, const float* near_jitter, int64_t N, const uint8_t* binaries, const float* aabbs_host,
                        int levels, int R, float step, float cone_angle, int max_samples, const int64_t* offsets, int32_t* ray_idx,
                        float* t0, float* t1, int32_t* cells, void* stream);
/* nerfacc.render_visibility_from_density: visible[i] = (T_i >= early_stop_eps) && (alpha_i >= alpha_thre) */
int acn_occ_visibility(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                       float early_stop_eps, float alpha_thre, uint8_t* visible, void* stream);
/* _merge_segments_union (ray_rendering.py:196-256) as a per-ray K-way merge.  out_offsets NULL = count pass (out_counts[r] =
 * merged segments of ray r); else write pass. */
typedef struct {
  const int64_t* offsets[ACN_MAX_EXPERTS];
  const int32_t* counts[ACN_MAX_EXPERTS];
  const float* t0[ACN_MAX_EXPERTS];
  const float* t1[ACN_MAX_EXPERTS];
  int32_t K;
} AcnSegLists;
int acn_seg_union(const AcnSegLists* lists, int64_t N, const int64_t* out_offsets, int32_t* out_counts, int32_t* ray_idx, float* t0,
                  float* t1, void* stream);
/* weights (M,) = alpha_i * prod_{j<i}(1 - alpha_j), alpha = 1 - exp(-sigma (t1 - t0)); rgb_map (N,3) = sum w rgb, depth (N,) =
 * sum w (t0+t1)/2, acc (N,) = sum w.  Backward: gradients w.r.t. sigma (M,) and rgb (M,3); any d_* may be NULL. */
int acn_packed_render_fwd(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                          const float* rgb, float* weights, float* rgb_map, float* depth, float* acc, void* stream);
int acn_packed_render_bwd(const int64_t* offsets, const int32_t* counts, int64_t N, const float* t0, const float* t1, const float* sigma,
                          const float* rgb, const float* weights, const float* d_rgb_map, const float* d_depth, const float* d_acc,
                          const float* d_weights, float* g_sigma, float* g_rgb, void* stream);

/* ---- (f)-3: episode-sampler routing (data/task_dataset.py:130-171, 242-351, 353-418, 544-627) --------------------------
 * Per ray: clip to the region box and [near, far] (valid), pick the micro-cell (policy 0 = "dda": the cell with the longest
 * in-cell path along a 64-step grid walk; policy 1 = "alpha": the cell of the alpha point or the best of its 6 neighbours by
 * overlap), and test the overlap with the chosen cell against that cell's tolerance (keep).  rays (N, ncol) with ncol = 6 or
 * >= 8 ([o, d, near, far, ...]); aabb_host [lo(3), hi(3)]; cells_host (nx, ny, nz); cell_bounds (C, 6) [lo, hi] per cell and
 * tol_per_cell (C,) on the device (TaskDataset.cell_bounds and max(1e-6 |cell diagonal|, 1e-9)); tol_abs: the alpha policy's
 * absolute tolerance.  cid = -1 for invalid rays.  Integer outputs are bit-exact against the reference. */
int acn_episode_route(const float* rays, int ncol, int64_t N, const float* aabb_host, const int32_t* cells_host,
                      const float* cell_bounds, const float* tol_per_cell, int policy, float alpha, float tol_abs, int max_steps,
                      uint8_t* valid, int32_t* cid, float* len_best, uint8_t* keep, void* stream);

/* ---- (f)-4: Voronoi mask builder (scripts/create_clusters.py:386-634; reductions :898-904, 928-932) ----------------------
 * rays (N,8) [o, d, near, far]; per ray `ray_samples` positions at lerp(near, far, linspace(0,1,S)); a sample belongs to the
 * centroids with d^2 <= margin^2 * min d^2 (margin == 1: the nearest, first index on ties), distances over components (1,2)
 * when cluster_2d else (0,1,2); mask (N,C) uint8 = any sample of the ray belongs to c.  mins_io / maxs_io (C,3) fp32 and
 * counts_io (C,) int64 (all three or none): running per-expert bounding box and count of the member sample positions,
 * updated in place (the caller all-reduces them MIN / MAX / SUM across ranks); ws: acn_voronoi_ws_bytes() bytes. */
int64_t acn_voronoi_ws_bytes(void);
int acn_voronoi_masks(const float* rays, int64_t N, int ray_samples, const float* centroids_host, int C, int cluster_2d, float margin,
                      uint8_t* mask, float* mins_io, float* maxs_io, int64_t* counts_io, void* ws, void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* ACNERF_H_ */
