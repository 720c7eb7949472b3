/*
 * rpc_b200.h -- C-ABI of the B200 (sm_100a) hot path of temex12/RobustPointClouds:
 *   perturb (scale, clamp, add)  ->  hard / dynamic voxelization  ->  HardSimpleVFE mean
 *   -> backward to the per-point raw perturbation.
 *
 * Drop-in boundary.  Every entry point takes plain pointers and sizes, a CUDA stream
 * as `void*` (a `cudaStream_t`), returns an `int` status (0 = OK, <0 = error, see rpc_strerror),
 * never throws, never allocates, never synchronises the host, and launches a fixed topology of
 * kernels, so a sequence of calls is CUDA-graph capturable.  Pointers marked [dev] are device
 * memory owned by the caller for the duration of the stream-ordered call; pointers marked [host]
 * are read on the host before the call returns (their values are baked into kernel parameters).
 *
 * What each entry point replaces on the reference side:
 *   rpc_voxelize_hard      mmcv `_ext.hard_voxelize_forward` as driven by mmdet3d
 *                          `Det3DDataPreprocessor.voxelize` (whole batch in one call, batch index
 *                          already in coors[:,0]); consumed at
 *                          /root/reference/models/detectors/adversarial_voxelnet.py:66-71 and
 *                          /root/reference/models/detectors/adversarial_centerpoint.py:62,191
 *   rpc_voxelize_dynamic   mmcv `_ext.dynamic_voxelize_forward` (same call sites, dynamic mode)
 *   rpc_vfe_mean_fwd/bwd   mmdet3d `HardSimpleVFE.forward` and its autograd, called at
 *                          adversarial_voxelnet.py:135-137, adversarial_centerpoint.py:100-102
 *   rpc_valid_rows         the padding mask + boolean gather of adversarial_voxelnet.py:86-90
 *   rpc_scatter_rows / rpc_gather_rows   the masked scatter-back (adversarial_voxelnet.py:113-117) and its
 *                          autograd
 *   rpc_perturb_apply_fwd/bwd   VoxelPerturber's scale/clamp/add tail
 *                          (/root/reference/models/adversarial/voxel_perturber.py:256,357-365,285)
 *   rpc_voxelize_hard (raw != NULL) + rpc_step_bwd   the north-star order: perturb -> hard voxelize ->
 *                          VFE mean, and the gradient back to the raw per-point perturbation
 *   rpc_dynamic_scatter_mean_fwd   dynamic voxelization + mmcv DynamicScatter(mean) in one launch for the
 *                          dynamic configuration; its backward is rpc_step_bwd
 */
#ifndef RPC_B200_H_
#define RPC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RPC_ABI_VERSION 1
#define RPC_MAX_BATCH 32      /* sweeps per call */
#define RPC_MAX_CHANNELS 8    /* features per point, 3..8 */

enum {
    RPC_OK = 0,
    RPC_ERR_NULL = -1,        /* a required pointer is NULL */
    RPC_ERR_CONFIG = -2,      /* C, P, V, voxel size or range invalid */
    RPC_ERR_BATCH = -3,       /* B outside 1..RPC_MAX_BATCH or a negative point count */
    RPC_ERR_WORKSPACE = -4,   /* workspace smaller than rpc_workspace_bytes() */
    RPC_ERR_CUDA = -5,        /* a CUDA launch failed (cudaGetLastError) */
    RPC_ERR_TOO_LARGE = -6,   /* point count or grid does not fit 31-bit indices */
    RPC_ERR_ALIGN = -7,       /* a pointer is not 16-byte aligned */
    RPC_ERR_STATIC_LIMIT = -8 /* input beyond the fixed capacity of a *_static entry point */
};

/* mirrors mmcv Voxelization(voxel_size, point_cloud_range, max_num_points, max_voxels) */
typedef struct rpc_voxel_cfg {
    float voxel_size[3];      /* (vx, vy, vz) */
    float pcr[6];             /* (x_lo, y_lo, z_lo, x_hi, y_hi, z_hi) */
    int32_t max_points;       /* P, >= 1 (hard) */
    int32_t max_voxels;       /* V per sweep, >= 1 (hard) */
    int32_t C;                /* features per point, 3..RPC_MAX_CHANNELS */
    int32_t F;                /* HardSimpleVFE num_features, 1..C */
} rpc_voxel_cfg;

/* per-channel bounds of VoxelPerturber (voxel_perturber.py:209-256 eb, :333-359 fb) */
typedef struct rpc_perturb_cfg {
    float eb[RPC_MAX_CHANNELS];   /* p = raw * eb[c] */
    float fb[RPC_MAX_CHANNELS];   /* p = clamp(p, -fb[c], +fb[c]); NaN -> 0 */
} rpc_perturb_cfg;

int rpc_abi_version(void);
const char *rpc_strerror(int status);

/* Per-stage device timing of the MOST RECENT rpc_voxelize_hard / rpc_step_bwd call (bench.py's roofline
 * leg; no reference counterpart).  While enabled every kernel launch of a call is bracketed by
 * cudaEvents on the call's stream; do not enable while capturing a CUDA graph.  rpc_profile_read waits
 * for the call to finish and returns the number of stages, their durations (ms) and kernel names. */
void rpc_profile_enable(int on);
int rpc_profile_read(float *ms, const char **names, int max_stages);

/* grid[d] = round((hi-lo)/vs) in fp32, d = x,y,z (mmcv voxelization.cpp) */
int rpc_grid_size(const rpc_voxel_cfg *cfg, int32_t grid[3]);

/* bytes of scratch rpc_voxelize_hard / rpc_step_fwd / rpc_dynamic_scatter_mean_fwd need for a batch of B sweeps
 * with n_per_sample[b] points each; 0 on invalid arguments */
size_t rpc_workspace_bytes(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample /*[host]*/);
/* byte offset, inside the workspace, of 8 uint64 %globaltimer stamps (ns) that rpc_voxelize_hard writes at
 * its phase boundaries while profiling is enabled (rpc_profile_enable): start, after clear, insert, flag,
 * assign, rank, end of CTA 0's gather.  -1 on invalid arguments. */
int64_t rpc_workspace_stamps_offset(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample /*[host]*/);
/* rows of the voxel outputs the caller must provide: sum_b min(n_b, V) */
int64_t rpc_voxel_capacity(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample /*[host]*/);

/*
 * Hard voxelization of a batch of B sweeps (points of all sweeps concatenated, sweep b holding
 * n_per_sample[b] rows), optionally of the PERTURBED points x + clamp(raw*eb, +-fb) when `raw`
 * and `pc` are non-NULL.  Outputs are compact and concatenated in sweep order, exactly what
 * Det3DDataPreprocessor.voxelize returns after torch.cat:
 *   voxels      [cap, P, C] f32   rows >= voxel_total are not written; unused slots are 0.0
 *   coors       [cap, 4]   i32    (batch, z, y, x)
 *   num_points  [cap]      i32
 *   voxel_num   [B + 1]    i32    per-sweep voxel counts, then their total      (device memory)
 *   point2voxel [ntotal]   i32    global voxel row of each input point, -1 if not stored (may be NULL)
 *   point2slot  [ntotal]   i32    slot 0..P-1 of each input point, -1 if not stored (may be NULL)
 *   feats       [cap, F]   f32    HardSimpleVFE mean of the rows (may be NULL)
 * Voxel order = order of the first point of each cell; slot order = point order; first P kept;
 * cells first seen after V voxels exist are refused (mmcv hard_voxelize_forward_cpu semantics).
 */
int rpc_voxelize_hard(const rpc_voxel_cfg *cfg, const float *points /*[dev]*/,
                      const int32_t *n_per_sample /*[host]*/, int B,
                      const float *raw /*[dev] or NULL*/, const rpc_perturb_cfg *pc /*[host] or NULL*/,
                      float *voxels, int32_t *coors, int32_t *num_points, int32_t *voxel_num,
                      int32_t *point2voxel, int32_t *point2slot, float *feats,
                      void *workspace, size_t workspace_bytes, void *stream);

/*
 * Backward of the fused step (rpc_voxelize_hard with raw + feats), per input point i:
 *   g[c] = grad_feats[v, c] / num_points[v] when point i was stored in voxel v = point2voxel[i]
 *          (c < F), else 0
 *   grad_raw[i, c] = (-fb[c] <= raw[i,c]*eb[c] <= fb[c]) ? g[c] * eb[c] : 0
 * which is autograd through HardSimpleVFE (adversarial_voxelnet.py:135-137), the masked
 * scatter-back (:113-114) and VoxelPerturber's clamp and scale (voxel_perturber.py:256,357-359).
 * With raw == NULL the gradient w.r.t. the point features (g) is written instead.
 */
int rpc_step_bwd(const rpc_voxel_cfg *cfg, const float *grad_feats /*[cap,F]*/,
                 const int32_t *num_points, const int32_t *point2voxel, const float *raw,
                 const rpc_perturb_cfg *pc /*[host]*/, int64_t ntotal, float *grad_raw /*[ntotal,C]*/,
                 void *stream);

/* dynamic voxelization: coors [ntotal, ncol]; ncol == 4: (batch, z, y, x) rows as
 * Det3DDataPreprocessor.voxelize returns them, ncol == 3: (z, y, x) as the Voxelization module
 * returns them; the three cell columns are all -1 for a dropped point (mmcv CPU rule).
 * `raw`/`pc` as above. */
int rpc_voxelize_dynamic(const rpc_voxel_cfg *cfg, const float *points,
                         const int32_t *n_per_sample /*[host]*/, int B, const float *raw,
                         const rpc_perturb_cfg *pc, int32_t *coors, int ncol, void *stream);

/*
 * Dynamic voxelization + mmcv DynamicScatter(mean) of a batch (BASELINE.json config 4; mmdet3d
 * DynamicSimpleVFE / `voxel_type='dynamic'` consumers of the call sites above), one launch.
 *   coors_in    [ntotal, 4] i32 (b, z, y, x) or NULL.  NULL: cells are computed from the (perturbed, when
 *               raw/pc are given) points and sweep b owns rows n_per_sample-wise; non-NULL: points are
 *               scattered by these rows (DynamicScatter.forward(points, coors)); rows with a negative or
 *               out-of-grid entry are dropped.
 *   coors_out   [ntotal, 4] i32 per-point (b, z, y, x), cells -1 when dropped (may be NULL; ignored with coors_in)
 *   point2voxel [ntotal]    i32 voxel row of each point, -1 when dropped
 *   feats       [cap, F]    f32 mean of the (perturbed) points of the voxel, all points, no cap
 *   voxel_coors [cap, 4]    i32 (b, z, y, x); per sample ascending in (z, y, x) (torch.unique order), samples in order
 *   num_points  [cap]       i32 points per voxel
 *   voxel_num   [B + 1]     i32 per-sample voxel counts, then the total            (device memory)
 * cap = ntotal rows (every point its own voxel).  Sums are fp32 red.add (order-free; mmcv's CUDA reduce is
 * atomic too).  The backward is rpc_step_bwd(grad_feats, num_points, point2voxel, raw, ...).
 */
size_t rpc_dynamic_workspace_bytes(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample /*[host]*/);
/* byte offset, inside the workspace, of 10 uint64 %globaltimer stamps (ns) that CTA 0 of the dynamic kernel
 * writes at its phase boundaries while profiling is enabled: start, after clear, mark, coarse sums, coarse
 * scan, fine bits, fine sums, rank, sum, mean.  -1 on invalid arguments. */
int64_t rpc_dynamic_stamps_offset(const rpc_voxel_cfg *cfg, int B, const int32_t *n_per_sample /*[host]*/);
int rpc_dynamic_scatter_mean_fwd(const rpc_voxel_cfg *cfg, const float *points,
                                 const int32_t *n_per_sample /*[host]*/, int B, const float *raw,
                                 const rpc_perturb_cfg *pc, const int32_t *coors_in, int32_t *coors_out,
                                 int32_t *point2voxel, float *feats, int32_t *voxel_coors, int32_t *num_points,
                                 int32_t *voxel_num, void *workspace, size_t workspace_bytes, void *stream);

/* HardSimpleVFE: out[m, c] = sum_k voxels[m, k, c] / num_points[m], c < F, slots summed in order.
 * m_dev (device int32, may be NULL) bounds the rows at run time: only min(M, *m_dev) rows are done. */
int rpc_vfe_mean_fwd(const float *voxels, const int32_t *num_points, int64_t M, int P, int C, int F,
                     const int32_t *m_dev, float *out, void *stream);
/* its autograd: grad_voxels[m, k, c] = grad_out[m, c] / num_points[m] (every slot), 0 for c >= F */
int rpc_vfe_mean_bwd(const float *grad_out, const int32_t *num_points, int64_t M, int P, int C, int F,
                     const int32_t *m_dev, float *grad_voxels, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Reference-order composition: voxelize first (no perturbation), then perturb the kept points inside the
 * padded voxel tensor, as AdversarialVoxelNet.extract_feat / AdversarialCenterPoint.extract_pts_feat do.
 * ------------------------------------------------------------------------------------------------- */

/* scratch of rpc_valid_rows for an [M,P,*] tensor */
size_t rpc_valid_rows_workspace_bytes(int64_t M, int P);

/* extract_feat's padding mask + boolean gather (adversarial_voxelnet.py:86-90, adversarial_centerpoint.py:
 * 73-77) on the [M*P, C] view of `voxels`:  mode 0 'rowsum': valid = fp32 row sum != 0 (the reference's
 * rule, which also drops real points whose features sum to 0); mode 1 'count': valid = slot < num_points.
 *   row_idx [M*P] i32   k-th valid row, ascending (first count[0] entries defined)
 *   row2k   [M*P] i32   k of each row or -1                     (may be NULL)
 *   rows_out[M*P, C]    the valid rows themselves = voxels_2d[valid_mask]   (may be NULL)
 *   count   [1]   i32   K', on the device (no host sync)
 * m_dev as in rpc_vfe_mean_fwd. */
int rpc_valid_rows(const float *voxels, const int32_t *num_points, int64_t M, int P, int C, int mode,
                   const int32_t *m_dev, int32_t *row_idx, int32_t *row2k, float *rows_out, int32_t *count,
                   void *workspace, size_t workspace_bytes, void *stream);

/* the gradient-connected copy + masked scatter-back (adversarial_voxelnet.py:113-117):
 * out = voxels; out_2d[row_idx[k]] = rows[k]  (through row2k) */
int rpc_scatter_rows(const float *voxels, const int32_t *row2k, const float *rows, int64_t M, int P, int C,
                     float *out, void *stream);
/* out[k, :] = src_2d[row_idx[k], :], k < min(K, *k_dev): the boolean gather, and the autograd of the scatter */
int rpc_gather_rows(const float *src, const int32_t *row_idx, int64_t K, const int32_t *k_dev, int C,
                    float *out, void *stream);

/* VoxelPerturber's elementwise tail (voxel_perturber.py:256, 357-365, 285):
 *   p = nan_to_num(clamp(raw * eb[c], -fb[c], +fb[c]));  y = x + p      (p_out may be NULL) */
int rpc_perturb_apply_fwd(const float *x, const float *raw, const rpc_perturb_cfg *pc /*[host]*/, int64_t K,
                          const int32_t *k_dev, int C, float *y, float *p_out, void *stream);
/* its autograd: grad_raw = (-fb <= raw*eb <= fb) ? grad_p * eb : 0 */
int rpc_perturb_apply_bwd(const float *grad_p, const float *raw, const rpc_perturb_cfg *pc /*[host]*/,
                          int64_t K, const int32_t *k_dev, int C, float *grad_raw, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * One-shot all-reduce (sum) of the perturber's flat parameter-gradient buffer over NVLink peer memory:
 * the path's only exchange (the reference gets it from mmengine's DDP wrapper, custom_hook.py:24-28,
 * configs/adversarial/adversarial-centerpoint_voxel-nuscenes.py:110).  stage0 / stage1 / peer_flags are HOST
 * arrays of `world` device pointers valid on this GPU (peer-mapped symmetric buffers, index = rank): two
 * staging buffers of n floats (rounded up to 4) and rpc_allreduce_flag_bytes() of flags per rank, flags and
 * `counter` (2 x uint32, local) zero-initialised once before the first call on any rank.  Every rank calls it
 * the same number of times.  out[i] = sum over ranks of src[i], added in rank order (bit-identical on all
 * ranks); out may alias src.  The launch parameters do not change between calls: capturable in a CUDA graph.
 * *status (device int32) is set to 1 if a peer did not arrive within 2 s.
 * ------------------------------------------------------------------------------------------------- */
size_t rpc_allreduce_flag_bytes(void);
int rpc_allreduce_oneshot(float *const *stage0 /*[host]*/, float *const *stage1 /*[host]*/,
                          uint32_t *const *peer_flags /*[host]*/, int rank, int world, int64_t n, const float *src,
                          float *out, uint32_t *counter, int32_t *status, void *stream);

/* Low-latency variant of the same exchange: every rank pushes each value together with the call number as one
 * 8-byte word into every peer's receive area and polls its own -- one one-way NVLink latency per call, no flag
 * round trips.  recv[r] (host array of `world` device pointers) = rank r's peer-mapped receive area of
 * rpc_allreduce_ll_bytes(world, n) bytes, zeroed once before the first call; counter / status as above. */
size_t rpc_allreduce_ll_bytes(int world, int64_t n);
int rpc_allreduce_ll(void *const *recv /*[host]*/, int rank, int world, int64_t n, const float *src, float *out,
                     uint32_t *counter, int32_t *status, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Hand-off to the middle encoder.  coors rows are already (batch, z, y, x) int32, the order
 * spconv's SparseConvTensor takes (adversarial_voxelnet.py:140-141, adversarial_centerpoint.py:105-106), and
 * the batch size is the caller's B (no coors[-1, 0].item()).  rpc_voxel_centers adds the `voxel_centers` output of
 * mmdet3d Det3DDataPreprocessor.voxelize: centers[v] = (coors[v, (x, y, z)] + 0.5) * voxel_size + range_min,
 * fp32, multiply and add unfused like the ATen expression; coors [M, ncol] with ncol 3 (z, y, x) or 4; only the
 * first min(M, *m_dev) rows are written when m_dev (device) is given.
 * ------------------------------------------------------------------------------------------------- */
int rpc_voxel_centers(const rpc_voxel_cfg *cfg, const int32_t *coors, int ncol, int64_t M, const int32_t *m_dev,
                      float *centers, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * The elementwise / reduction parts of VoxelPerturber.forward around its MLP, without host
 * syncs.  `moments` is rpc_moments_bytes() of device scratch (fp64 accumulators, zeroed by the call).
 *   rpc_normalize_fwd       xn = clamp(x / (std(x, dim 0) + 1e-6), -10, 10); NaN / Inf scale -> unit scaling
 *                           (voxel_perturber.py:157-168)
 *   rpc_perturb_losses_fwd  losses[4] = {l2_norm = mean_k ||p_k||, intensity_loss = mean_k |p_k3| (0 for C < 4),
 *                           bias_loss = mean_c |mean_k p_kc|, imbalance_loss = std_c(std_k p_kc)}
 *                           (voxel_perturber.py:268-299); consts[32] f32 keeps what the backward needs
 *   rpc_perturb_losses_bwd  grad_p [K, C] from grad_losses[4] (device), torch's autograd rules (norm / std
 *                           backward are 0 where the norm / std is 0)
 * ------------------------------------------------------------------------------------------------- */
size_t rpc_moments_bytes(void);
int rpc_normalize_fwd(const float *x, int64_t K, const int32_t *k_dev, int C, double *moments, float *xn,
                      void *stream);
int rpc_perturb_losses_fwd(const float *p, int64_t K, const int32_t *k_dev, int C, double *moments, float *losses,
                           float *consts, void *stream);
int rpc_perturb_losses_bwd(const float *p, const float *consts, const float *grad_losses, int64_t K,
                           const int32_t *k_dev, int C, float *grad_p, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Pre-path pipeline op: mmdet3d PointsRangeFilter (configs/_base_/kitti-3d-car.py:62-64 of the
 * reference) = points[BasePoints.in_range_3d(range)], the OPEN box lo < p < hi on x, y, z (NaN dropped), order kept.
 * out_points [N, C] (first count[0] rows written; may be NULL), out_index [N] = source row of every kept point,
 * count[0] on the device; workspace of rpc_valid_rows_workspace_bytes(N, 1).  range6 is a HOST array.
 * ------------------------------------------------------------------------------------------------- */
int rpc_points_range_filter(const float *points, int64_t N, int C, const float *range6 /*[host]*/, float *out_points,
                            int32_t *out_index, int32_t *count, void *workspace, size_t workspace_bytes, void *stream);

/*
 * Allocation-free variant of rpc_points_range_filter.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight
 * serving loops, harnesses that forbid it) and therefore cannot provide a workspace.  Same inputs, outputs and
 * results as rpc_points_range_filter; the per-tile count scratch lives in a `static __device__` array inside the
 * library (16 640 bytes, held by every GPU that loads the library until the process exits), so
 *   - N is limited to RPC_STATIC_FILTER_MAX_POINTS rows; a larger N returns RPC_ERR_STATIC_LIMIT and launches nothing;
 *   - the call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine);
 *   - every call on a device shares the array: calls that overlap on the GPU (different streams, or graphs
 *     replayed concurrently) race on it; stream-ordered calls are safe.  The kernel writes every count before it
 *     reads it, so stale contents of the previous call are harmless.
 */
#define RPC_STATIC_FILTER_MAX_POINTS 4194304   /* 4096 tiles of 1024 rows */
int rpc_points_range_filter_static(const float *points, int64_t N, int C, const float *range6 /*[host]*/,
                                   float *out_points, int32_t *out_index, int32_t *count, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Post-VFE perturbation variant: the elementwise part of
 * StrongAdversarialVoxelNet.apply_enhanced_perturbations (models/detectors/strong_adversarial_voxelnet.py:156-180)
 * on n = M * C features: sp = (y - x) * scaling [+ alpha * last] (last may be NULL: no momentum yet; y_is_delta != 0:
 * the adversary returned the perturbation itself, :159-161, and y - x is replaced by y);
 * out = x + sp; sp_out = sp (the next call's `last`); l2[0] = ||sp||_2 (device; sumsq: 8 bytes of device scratch).
 * The backward takes the gradients on out (g_out, may be NULL) and on l2 (g_l2, device scalar, may be NULL) and
 * writes grad_x and / or grad_y (either may be NULL).
 * ------------------------------------------------------------------------------------------------- */
int rpc_momentum_apply_fwd(const float *x, const float *y, const float *last, int y_is_delta, float scaling,
                           float alpha, int64_t n, float *out, float *sp_out, double *sumsq, float *l2, void *stream);
int rpc_momentum_apply_bwd(const float *g_out, const float *g_l2, const float *sp, const float *l2, int y_is_delta,
                           float scaling,
                           int64_t n, float *grad_x, float *grad_y, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Spatial-attention gate of VoxelPerturber (first slice of SURVEY 8f row 2):
 *   out[k, :] = t[k, :] * a[k],  a = sigmoid(w2 . relu(W1 xn[k, :] + b1) + b2)
 * = `raw_perturbations * self.attention(voxel_features_normalized)` (models/adversarial/voxel_perturber.py:203-205)
 * with attention = Linear(C, mid) -> ReLU -> Linear(mid, 1) -> Sigmoid, mid = max(C / 2, 1) (:104-110).
 * xn, t, out [K, C] f32; w1 [mid, C], b1 [mid], w2 [1, mid], b2 [1] are the nn.Linear parameter tensors as they lie;
 * gate [K] (the a values) may be NULL.  One streaming pass instead of five ATen launches.
 * The backward recomputes the gate (nothing of size K is saved): grad_t [K, C] = g_out * a, grad_xn [K, C] (may be
 * NULL), and the parameter gradients, reduced in fp64 through `scratch` (rpc_attention_scratch_bytes() of device
 * memory, zeroed by the call) into grad_w1 / grad_b1 / grad_w2 / grad_b2 (shapes of the parameters, overwritten).
 * ------------------------------------------------------------------------------------------------- */
size_t rpc_attention_scratch_bytes(void);
int rpc_attention_gate_fwd(const float *xn, const float *t, const float *w1, const float *b1, const float *w2,
                           const float *b2, int64_t K, int C, float *out, float *gate, void *stream);
int rpc_attention_gate_bwd(const float *g_out, const float *xn, const float *t, const float *w1, const float *b1,
                           const float *w2, const float *b2, int64_t K, int C, double *scratch, float *grad_t,
                           float *grad_xn, float *grad_w1, float *grad_b1, float *grad_w2, float *grad_b2,
                           void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Weight and bias gradient of one Linear layer of the perturber's MLP (second slice of SURVEY 8f row 2;
 * models/adversarial/voxel_perturber.py:82-101 builds the layers, autograd's addmm backward is what the reference runs):
 *   grad_w[m, n] = sum_k g[k, m] * a[k, n]   (nn.Linear weight layout [out = M, in = N]),  grad_b[m] = sum_k g[k, m]
 * g [K, M] = gradient of the layer's output, a [K, N] = the layer's input, both row-major f32; M, N <= 128.
 * grad_w [M, N] and grad_b [M] (may be NULL) are overwritten.  fp32 accumulation, split over the rows.
 * ------------------------------------------------------------------------------------------------- */
int rpc_linear_wgrad(const float *g, const float *a, int64_t K, int M, int N, float *grad_w, float *grad_b,
                     void *stream);
/* the same with the layer input given as the previous layer's pre-activation: a = relu(batch_norm(a_pre)) with the saved
 * statistics bn_mean / bn_invstd / bn_gamma / bn_beta [N], recomputed while the operand is staged (all four NULL: a_pre is
 * the input itself) */
int rpc_linear_wgrad_bn(const float *g, const float *a_pre, const float *bn_mean, const float *bn_invstd,
                        const float *bn_gamma, const float *bn_beta, int64_t K, int M, int N, float *grad_w, float *grad_b,
                        void *stream);

/* ---------------------------------------------------------------------------------------------------
 * Training-mode BatchNorm1d + ReLU of the perturber's MLP (third slice of SURVEY 8f row 2;
 * `nn.BatchNorm1d(h), nn.ReLU(inplace=True)` of models/adversarial/voxel_perturber.py:84-98, eps / momentum as set at :461-462):
 *   z = relu((y - mean) * invstd * gamma + beta), mean / biased variance over the K rows of y [K, H], H <= 128;
 *   save_mean / save_invstd [H] for the backward; running_mean / running_var [H] (may be NULL) updated like torch
 *   (momentum, unbiased variance).  `scratch` = rpc_bn_scratch_bytes() of device memory (fp64 sums, zeroed by the call).
 * Backward: g = dL/dz; the ReLU mask is recomputed from y; grad_y [K, H], grad_gamma / grad_beta [H] (may be NULL).
 * ------------------------------------------------------------------------------------------------- */
size_t rpc_bn_scratch_bytes(void);
int rpc_bn_relu_fwd(const float *y, int64_t K, int H, const float *gamma, const float *beta, float eps, float momentum,
                    float *running_mean, float *running_var, double *scratch, float *save_mean, float *save_invstd,
                    float *z, void *stream);
int rpc_bn_relu_bwd(const float *g, const float *y, const float *save_mean, const float *save_invstd,
                    const float *gamma, const float *beta, int64_t K, int H, double *scratch, float *grad_y,
                    float *grad_gamma, float *grad_beta, void *stream);

/* ---------------------------------------------------------------------------------------------------
 * The Linear layers of the perturber's MLP trunk on the tensor cores (fourth slice of SURVEY 8f row 2): the forward
 * products of `self.model` (models/adversarial/voxel_perturber.py:82-103, run at :176) and the input-gradient products
 * autograd's addmm backward derives from them -- what the reference gets from cuBLAS.
 *   y[K, NO] = f(x)[K, NI] * B^T (+ bias) (-> tanh),  NI, NO <= 128
 *   w_layout 0: B = w [NO, NI] (forward, the nn.Linear weight as it lies); 1: B^T = w [NI, NO] (input gradient: x = dL/dy
 *   of the layer, w = the same weight tensor)
 *   in_mean / in_invstd / in_gamma / in_beta [NI] (all or none): f = the previous layer's training-mode BatchNorm +
 *   ReLU applied while x is loaded, so the normalised activation is never written; epi_tanh: the trunk's final Tanh (:102)
 *   moments (fp64 [2 * NO], may be NULL; zeroed by the call): += column sums and sums of squares of y, the statistics
 *   of the BatchNorm that follows; rpc_bn_finalize turns them into save_mean / save_invstd (biased variance, eps) and
 *   updates the running statistics like torch; rpc_bn_relu_apply recomputes z = relu(bn(y)) from saved statistics
 *   (the layer input the weight gradient needs in the backward).
 * NI > 8 (NI % 8 == 0, any NO): 3 x TF32 on tcgen05 (x = x_hi + x_lo, w = w_hi + w_lo, x_lo w_hi + x_hi w_lo + x_hi w_hi
 * accumulated in fp32 in tensor memory; ~1e-6 relative; the weights stay in 64 KB of shared memory, which rules out 128 -> 128 and 128 <-> 96: RPC_ERR_CONFIG).  NI <= 8 (first layer, input gradient of the last; NO % 4 == 0): an
 * fp32 FMA stream.
 * k_dev (device int32, may be NULL) bounds the rows at run time.
 * ------------------------------------------------------------------------------------------------- */
int rpc_linear_fwd(const float *x, int64_t K, const int32_t *k_dev, int NI, int NO, const float *w, int w_layout,
                   const float *bias, const float *in_mean, const float *in_invstd, const float *in_gamma,
                   const float *in_beta, int epi_tanh, double *moments, float *y, void *stream);
int rpc_bn_finalize(const double *moments, int64_t K, const int32_t *k_dev, int H, float eps, float momentum,
                    float *running_mean, float *running_var, float *save_mean, float *save_invstd, void *stream);
int rpc_bn_relu_apply(const float *y, int64_t K, const int32_t *k_dev, int H, const float *mean, const float *invstd,
                      const float *gamma, const float *beta, float *z, void *stream);
/* Input gradient of a Linear layer (grad_in [K, NO] = g [K, NI] * w, w [NI, NO] = the nn.Linear weight) whose epilogue also
 * reduces the backward of the BatchNorm + ReLU BELOW it: v [K, NO] is that layer's pre-activation with statistics v_mean /
 * v_invstd / v_gamma / v_beta, and red (fp64 [2 * NO], zeroed by the call) receives sum g' and sum g' * vhat with g' =
 * grad_in where relu(bn(v)) was open -- what rpc_bn_relu_bwd's first pass computes.  rpc_bn_relu_bwd_sums is
 * rpc_bn_relu_bwd with those sums given (its reduction pass over g and y is skipped).  NI > 8 (tensor-core kernel only). */
int rpc_linear_dgrad_bnred(const float *g, int64_t K, int NI, int NO, const float *w, const float *v, const float *v_mean,
                           const float *v_invstd, const float *v_gamma, const float *v_beta, double *red, float *grad_in,
                           void *stream);
int rpc_bn_relu_bwd_sums(const float *g, const float *y, const float *save_mean, const float *save_invstd, const float *gamma,
                         const float *beta, int64_t K, int H, const double *sums, float *grad_y, float *grad_gamma,
                         float *grad_beta, void *stream);

/* The whole training-mode trunk in one call each way (the same launches as the per-layer entry points, issued from C):
 * `self.model(voxel_features_normalized)` of models/adversarial/voxel_perturber.py:176 and its autograd.
 * nl hidden layers Linear(width[l-1] -> width[l]) + BatchNorm1d + ReLU (width[-1] = C), then Linear(width[nl-1] -> C) + Tanh.
 *   y[l]   [K, width[l]]  pre-activations, kept for the backward (host array of nl device pointers)
 *   stats  [2 * sum width] per layer (mean [width], invstd [width]);  scratch: 2 * sum width doubles (forward),
 *          2 * 128 doubles (backward);  raw [K, C] = the trunk's output
 *   work   2 * K * max(width, C) floats of scratch for the backward
 *   grad_w / grad_b [nl + 1], grad_gamma / grad_beta [nl]: host arrays of device pointers, shapes of the parameters
 *   grad_xn [K, C] or NULL (the normalised input carries no gradient in extract_feat)
 * Widths: multiples of 8 in 8 .. 128, the last hidden one >= 16; 2 <= K. */
#define RPC_TRUNK_MAX_HIDDEN 6
typedef struct rpc_trunk_desc {
    int32_t C, nl;
    int32_t width[RPC_TRUNK_MAX_HIDDEN];
    const float *w[RPC_TRUNK_MAX_HIDDEN + 1], *b[RPC_TRUNK_MAX_HIDDEN + 1];      /* nl + 1 Linear layers */
    const float *gamma[RPC_TRUNK_MAX_HIDDEN], *beta[RPC_TRUNK_MAX_HIDDEN];
    float *running_mean[RPC_TRUNK_MAX_HIDDEN], *running_var[RPC_TRUNK_MAX_HIDDEN];    /* may be NULL */
    float eps[RPC_TRUNK_MAX_HIDDEN], momentum[RPC_TRUNK_MAX_HIDDEN];
} rpc_trunk_desc;
int rpc_trunk_fwd(const rpc_trunk_desc *t, const float *xn, int64_t K, float *const *y, float *stats, double *scratch,
                  float *raw, void *stream);
/* eval mode (no gradient): BatchNorm uses the running statistics, given per layer as (mean [width], invstd [width]) in
 * `stats`; work = 2 * K * max(width) floats; nothing is kept */
int rpc_trunk_fwd_eval(const rpc_trunk_desc *t, const float *xn, int64_t K, const float *stats, float *work, float *raw,
                       void *stream);
int rpc_trunk_bwd(const rpc_trunk_desc *t, const float *xn, int64_t K, float *const *y, const float *stats, const float *raw,
                  const float *g_raw, float *work, double *scratch, float *const *grad_w, float *const *grad_b,
                  float *const *grad_gamma, float *const *grad_beta, float *grad_xn, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RPC_B200_H_ */
