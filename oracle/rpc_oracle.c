/*
 * rpc_oracle.c -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * Plain-C restatement of the reference's per-step adversarial input path:
 *   perturb (scale, clamp, add)  ->  hard / dynamic voxelization  ->  HardSimpleVFE mean
 *   -> backward to the per-point raw perturbation.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file's shared object.  The product (robustpointclouds_b200/) never does.
 *
 * PARITY STATUS
 *   - perturb / mask / scatter-back / VFE / backward  (a-U4): PINNED.
 *     The reference's own VoxelPerturber class is imported from /root/reference by
 *     tests/golden/make_golden.py and its outputs are committed under tests/golden/; this file is
 *     checked against them in tests/test_oracle.py.
 *   - hard / dynamic voxelization (rows a-U2, a-U3) and DynamicScatter mean: PARITY UNPINNED.
 *     The arithmetic lives in mmcv (requirements.txt:9 "mmcv>=2.0.0", un-pinned, absent from
 *     /root/reference and from this image) and mmdet3d @ 962f093736ffe55c089bc618842a8b8567318c8c
 *     (.SUBMODULES.json:10-12, status "missing").  The reference has no tests / golden vectors
 *     (SURVEY 4).  What follows restates mmcv's published CPU algorithm
 *     (mmcv/ops/csrc/pytorch/cpu/voxelization.cpp: dynamic_voxelize_forward_cpu_kernel,
 *     hard_voxelize_forward_cpu_kernel; mmcv/ops/csrc/pytorch/cuda/scatter_points_cuda.cu:
 *     DynamicPointToVoxelForwardCUDAKernelLauncher), anchored on the reference's call sites:
 *       models/detectors/adversarial_voxelnet.py:66-71,135-137
 *       models/detectors/adversarial_centerpoint.py:62,100-102
 *       configs/adversarial/adversarial-second_hv_secfpn_8xb6-80e_kitti-3d-car.py:48-53
 *
 * Build: see oracle/Makefile  (gcc -O2 -ffp-contract=off: every fp32 op below is a single IEEE
 * operation, no FMA contraction -- a 1-ulp change flips a voxel at a cell boundary).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------- */
/* grid size: mmcv voxelization.cpp  grid_size[i] = round((range[3+i]-range[i]) / voxel_size[i]) */
/* evaluated in fp32 (std::vector<float>).                                                      */
ORC_API void orc_grid_size(const float vs[3], const float pcr[6], int grid[3])
{
    for (int d = 0; d < 3; ++d) {
        float g = (pcr[3 + d] - pcr[d]) / vs[d];
        grid[d] = (int)roundf(g);
    }
}

/* One point -> (z,y,x) integer cell, or failure.  mmcv dynamic_voxelize_forward_cpu_kernel:
 *     c = floor((points[i][j] - coors_range[j]) / voxel_size[j]);
 *     if (c < 0 || c >= grid_size[j]) { failed = true; break; }
 *     coor[ndim_minus_1 - j] = c;
 * `c` is an int: on x86 a NaN / out-of-int-range float converts to INT_MIN, i.e. "failed".  We
 * make that explicit so the result does not depend on UB (SURVEY 8a-U NaN note: target = drop). */
static inline int point_to_cell(const float *p, const float vs[3], const float pcr[6],
                                const int grid[3], int zyx[3])
{
    for (int j = 0; j < 3; ++j) {
        float f = floorf((p[j] - pcr[j]) / vs[j]);
        if (!(f >= 0.0f) || !(f < (float)grid[j])) return 0; /* NaN fails both */
        zyx[2 - j] = (int)f;
    }
    return 1;
}

/* dynamic voxelization: coors[N,3] (z,y,x), all three -1 when the point is dropped (CPU rule). */
ORC_API void orc_dynamic_voxelize(const float *points, int N, int C, const float vs[3],
                                  const float pcr[6], int *coors)
{
    int grid[3];
    orc_grid_size(vs, pcr, grid);
    for (int i = 0; i < N; ++i) {
        int zyx[3];
        if (point_to_cell(points + (size_t)i * C, vs, pcr, grid, zyx)) {
            coors[3 * i + 0] = zyx[0]; coors[3 * i + 1] = zyx[1]; coors[3 * i + 2] = zyx[2];
        } else {
            coors[3 * i + 0] = coors[3 * i + 1] = coors[3 * i + 2] = -1;
        }
    }
}

/* ---- sparse cell -> voxel index map (same results as the dense table, used for big grids in
 * tests so that a call does not have to fill 360 MB) ---------------------------------------- */
typedef struct { int64_t *keys; int *vals; size_t cap; } cellmap_t;
static int cellmap_init(cellmap_t *m, size_t n)
{
    size_t cap = 64; while (cap < 2 * n + 2) cap <<= 1;
    m->cap = cap;
    m->keys = (int64_t *)malloc(cap * sizeof(int64_t));
    m->vals = (int *)malloc(cap * sizeof(int));
    if (!m->keys || !m->vals) return -1;
    for (size_t i = 0; i < cap; ++i) m->keys[i] = -1;
    return 0;
}
static void cellmap_free(cellmap_t *m) { free(m->keys); free(m->vals); }
static inline int *cellmap_slot(cellmap_t *m, int64_t key)
{
    uint64_t h = (uint64_t)key * 0x9E3779B97F4A7C15ull;
    size_t s = (size_t)(h >> 20) & (m->cap - 1);
    for (;;) {
        if (m->keys[s] == key) return &m->vals[s];
        if (m->keys[s] == -1) { m->keys[s] = key; m->vals[s] = -1; return &m->vals[s]; }
        s = (s + 1) & (m->cap - 1);
    }
}

/* hard voxelization, mmcv hard_voxelize_forward_cpu_kernel semantics:
 *   - walk points in input order; dropped points skipped;
 *   - voxel index = order of FIRST point of each cell; once voxel_num == V new cells are refused
 *     (for good) but points of existing voxels are still accepted;
 *   - first P points of a voxel (in point order) fill slots 0..P-1; other slots stay 0.0;
 * table_mode 0: dense int32[gz*gy*gx] allocated and filled with -1 per call, exactly like
 *               mmcv's `coor_to_voxelidx = -at::ones({gz,gy,gx})`  (this is the timed CPU baseline);
 * table_mode 1: sparse map, identical results.
 * Outputs sized at V like mmcv (voxels zero-filled here); optional point2voxel/point2slot [N]
 * (-1 when the point is not stored).  Returns voxel_num, or <0 on allocation failure. */
ORC_API int orc_hard_voxelize(const float *points, int N, int C, const float vs[3],
                              const float pcr[6], int P, int V, int table_mode, float *voxels,
                              int *coors, int *num_points, int *point2voxel, int *point2slot)
{
    int grid[3];
    orc_grid_size(vs, pcr, grid);
    const int gx = grid[0], gy = grid[1], gz = grid[2];
    memset(voxels, 0, (size_t)V * P * C * sizeof(float));
    memset(coors, 0, (size_t)V * 3 * sizeof(int));
    memset(num_points, 0, (size_t)V * sizeof(int));

    int *dense = NULL; cellmap_t map = {0, 0, 0}; int have_map = 0;
    if (table_mode == 0) {
        size_t G = (size_t)gx * gy * gz;
        dense = (int *)malloc((G ? G : 1) * sizeof(int));
        if (!dense) return -1;
        for (size_t i = 0; i < G; ++i) dense[i] = -1;
    } else {
        if (cellmap_init(&map, (size_t)N)) return -1;
        have_map = 1;
    }

    int voxel_num = 0;
    for (int i = 0; i < N; ++i) {
        if (point2voxel) point2voxel[i] = -1;
        if (point2slot) point2slot[i] = -1;
        const float *p = points + (size_t)i * C;
        int zyx[3];
        if (!point_to_cell(p, vs, pcr, grid, zyx)) continue;
        int64_t cell = ((int64_t)zyx[0] * gy + zyx[1]) * gx + zyx[2];
        int *slot = dense ? &dense[cell] : cellmap_slot(&map, cell);
        int v = *slot;
        if (v == -1) {
            if (V != -1 && voxel_num >= V) continue;
            v = voxel_num++;
            *slot = v;
            coors[3 * v + 0] = zyx[0]; coors[3 * v + 1] = zyx[1]; coors[3 * v + 2] = zyx[2];
        }
        int n = num_points[v];
        if (P == -1 || n < P) {
            memcpy(voxels + ((size_t)v * P + n) * C, p, (size_t)C * sizeof(float));
            num_points[v] = n + 1;
            if (point2voxel) point2voxel[i] = v;
            if (point2slot) point2slot[i] = n;
        }
    }
    if (dense) free(dense);
    if (have_map) cellmap_free(&map);
    return voxel_num;
}

/* HardSimpleVFE.forward (mmdet3d voxel_encoders/voxel_encoder.py; call site
 * adversarial_voxelnet.py:135-137):
 *     features[:, :, :F].sum(dim=1) / num_points.type_as(features).view(-1, 1)
 * sums ALL P slots (padding rows contribute whatever they hold), sequentially in slot order. */
ORC_API void orc_vfe_mean_fwd(const float *voxels, const int *num_points, int M, int P, int C,
                              int F, float *out)
{
    for (int m = 0; m < M; ++m) {
        float n = (float)num_points[m];
        for (int c = 0; c < F; ++c) {
            float s = 0.0f;
            for (int k = 0; k < P; ++k) s += voxels[((size_t)m * P + k) * C + c];
            out[(size_t)m * F + c] = s / n;
        }
    }
}

/* autograd of the above: every slot (incl. padding) receives g[m,c] / n[m]; channels >= F get 0 */
ORC_API void orc_vfe_mean_bwd(const float *grad_out, const int *num_points, int M, int P, int C,
                              int F, float *grad_voxels)
{
    for (int m = 0; m < M; ++m) {
        float n = (float)num_points[m];
        for (int k = 0; k < P; ++k)
            for (int c = 0; c < C; ++c)
                grad_voxels[((size_t)m * P + k) * C + c] =
                    (c < F) ? grad_out[(size_t)m * F + c] / n : 0.0f;
    }
}

/* extract_feat's padding mask (adversarial_voxelnet.py:86-90, adversarial_centerpoint.py:73-77):
 *     valid_mask = voxels_2d.sum(dim=1) != 0          (mode 0, 'rowsum' -- the reference quirk)
 * or the slot < num_points rule                        (mode 1, 'count').
 * Row sum evaluated left to right in fp32.  Writes mask[rows]; returns the number of valid rows. */
ORC_API int orc_valid_mask(const float *voxels, const int *num_points, int M, int P, int C,
                           int mode, uint8_t *mask)
{
    int K = 0;
    for (int m = 0; m < M; ++m)
        for (int k = 0; k < P; ++k) {
            const float *r = voxels + ((size_t)m * P + k) * C;
            int ok;
            if (mode == 0) { float s = r[0]; for (int c = 1; c < C; ++c) s += r[c]; ok = (s != 0.0f); }
            else ok = (k < num_points[m]);
            mask[(size_t)m * P + k] = (uint8_t)ok;
            K += ok;
        }
    return K;
}

/* VoxelPerturber's elementwise tail (voxel_perturber.py:256, 357-359, 362-365, 285):
 *     p = raw * eb[c];  p = clamp(p, -fb[c], +fb[c]);  p = nan_to_num(p, 0);  y = x + p
 * torch.clamp(x,min,max) == min(max(x,min),max) with NaN propagated. */
static inline float clampf_torch(float v, float lo, float hi)
{
    if (v != v) return v;
    float t = v < lo ? lo : v;
    return t > hi ? hi : t;
}
ORC_API void orc_perturb_apply(const float *x, const float *raw, const float *eb, const float *fb,
                               int K, int C, float *y, float *p_out)
{
    for (size_t k = 0; k < (size_t)K; ++k)
        for (int c = 0; c < C; ++c) {
            float p = raw[k * C + c] * eb[c];
            p = clampf_torch(p, -fb[c], fb[c]);
            if (p != p) p = 0.0f;
            if (p_out) p_out[k * C + c] = p;
            y[k * C + c] = x[k * C + c] + p;
        }
}
/* d/d raw of the above: clamp passes the gradient where -fb <= raw*eb <= fb (inclusive, SURVEY
 * a-3 [PROBE]); NaN positions get 0. */
ORC_API void orc_perturb_bwd(const float *grad_p, const float *raw, const float *eb,
                             const float *fb, int K, int C, float *grad_raw)
{
    for (size_t k = 0; k < (size_t)K; ++k)
        for (int c = 0; c < C; ++c) {
            float p = raw[k * C + c] * eb[c];
            int pass = (p >= -fb[c]) && (p <= fb[c]);
            grad_raw[k * C + c] = pass ? grad_p[k * C + c] * eb[c] : 0.0f;
        }
}

/* The four regularisers of voxel_perturber.py:268-299, accumulated in double (arbiter):
 *   out[0] l2_norm   = mean_k ||p_k||_2
 *   out[1] intensity = mean_k |p[k,3]|           (0 when C < 4)
 *   out[2] bias      = mean_c | mean_k p[k,c] |
 *   out[3] imbalance = std_c( std_k p[k,c] )     (both unbiased, torch default)            */
ORC_API void orc_perturb_losses(const float *p, int K, int C, double out[4])
{
    double l2 = 0, inten = 0, sum[16] = {0}, sq[16] = {0};
    for (size_t k = 0; k < (size_t)K; ++k) {
        double n2 = 0;
        for (int c = 0; c < C; ++c) {
            double v = p[k * C + c];
            n2 += v * v; sum[c] += v;
        }
        l2 += sqrt(n2);
        if (C >= 4) inten += fabs((double)p[k * C + 3]);
    }
    double bias = 0, sd[16], sdm = 0;
    for (int c = 0; c < C; ++c) {
        double mean = sum[c] / K;
        for (size_t k = 0; k < (size_t)K; ++k) { double d = p[k * C + c] - mean; sq[c] += d * d; }
        bias += fabs(mean);
        sd[c] = sqrt(sq[c] / (K - 1));
        sdm += sd[c];
    }
    sdm /= C;
    double v = 0;
    for (int c = 0; c < C; ++c) v += (sd[c] - sdm) * (sd[c] - sdm);
    out[0] = l2 / K; out[1] = (C >= 4) ? inten / K : 0.0; out[2] = bias / C;
    out[3] = sqrt(v / (C - 1));
}

/* DynamicScatter(mean) (mmcv scatter_points: unique_dim(sorted) over (z,y,x) rows of ONE sample,
 * rows with a negative coord removed; feature = mean over all points of the voxel, no cap).
 * coors: [N,3].  Outputs: out_coors[M,3] sorted lexicographically, feats[M,C], map[N] (-1 if
 * dropped), count[M].  Sum accumulated in point order in fp32.  Returns M, <0 on alloc failure. */
typedef struct { int z, y, x, i; } dyn_key_t;
static int dyn_cmp(const void *a, const void *b)
{
    const dyn_key_t *p = (const dyn_key_t *)a, *q = (const dyn_key_t *)b;
    if (p->z != q->z) return p->z < q->z ? -1 : 1;
    if (p->y != q->y) return p->y < q->y ? -1 : 1;
    if (p->x != q->x) return p->x < q->x ? -1 : 1;
    return p->i < q->i ? -1 : (p->i > q->i);
}
ORC_API int orc_dynamic_scatter_mean(const float *points, const int *coors, int N, int C,
                                     float *feats, int *out_coors, int *map, int *count)
{
    dyn_key_t *keys = (dyn_key_t *)malloc((size_t)(N ? N : 1) * sizeof(dyn_key_t));
    if (!keys) return -1;
    int n = 0;
    for (int i = 0; i < N; ++i) {
        map[i] = -1;
        const int *c = coors + 3 * (size_t)i;
        if (c[0] < 0 || c[1] < 0 || c[2] < 0) continue;
        keys[n].z = c[0]; keys[n].y = c[1]; keys[n].x = c[2]; keys[n].i = i; ++n;
    }
    qsort(keys, (size_t)n, sizeof(dyn_key_t), dyn_cmp);
    int M = 0;
    for (int k = 0; k < n; ++k) {
        if (k == 0 || keys[k].z != keys[k - 1].z || keys[k].y != keys[k - 1].y ||
            keys[k].x != keys[k - 1].x) {
            out_coors[3 * M] = keys[k].z; out_coors[3 * M + 1] = keys[k].y;
            out_coors[3 * M + 2] = keys[k].x; count[M] = 0;
            for (int c = 0; c < C; ++c) feats[(size_t)M * C + c] = 0.0f;
            ++M;
        }
        map[keys[k].i] = M - 1;
        count[M - 1] += 1;
    }
    free(keys);
    for (int i = 0; i < N; ++i) {
        int m = map[i];
        if (m < 0) continue;
        for (int c = 0; c < C; ++c) feats[(size_t)m * C + c] += points[(size_t)i * C + c];
    }
    for (int m = 0; m < M; ++m)
        for (int c = 0; c < C; ++c) feats[(size_t)m * C + c] /= (float)count[m];
    return M;
}
/* its autograd rule: grad_point[i] = grad_feats[map[i]] / count[map[i]]  (0 when dropped) */
ORC_API void orc_dynamic_scatter_mean_bwd(const float *grad_feats, const int *map,
                                          const int *count, int N, int C, float *grad_points)
{
    for (int i = 0; i < N; ++i) {
        int m = map[i];
        for (int c = 0; c < C; ++c)
            grad_points[(size_t)i * C + c] =
                m < 0 ? 0.0f : grad_feats[(size_t)m * C + c] / (float)count[m];
    }
}

/* ------------------------------------------------------------------------------------------- */
/* Whole-step CPU baseline, ONE sweep, north-star order (perturb -> hard voxelize -> VFE mean ->
 * backward to the raw perturbation).  Used by bench.py (cpu_baseline / --impl reference) so the
 * timed CPU work is one C call per sweep; tests compose the primitives themselves.
 *   points[N,C], raw[N,C] -> feats[V,F] (first voxel_num rows valid), grad_raw[N,C]
 *   grad_feats[V,F] is the synthetic upstream gradient.
 * scratch buffers are caller-provided so that timing excludes malloc of the API outputs but
 * INCLUDES the dense table allocate+fill (table_mode 0), as mmcv does per call.              */
ORC_API int orc_step_fused(const float *points, const float *raw, int N, int C, const float *eb,
                           const float *fb, const float vs[3], const float pcr[6], int P, int V,
                           int F, int table_mode, const float *grad_feats,
                           float *perturbed /*[N,C]*/, float *voxels /*[V,P,C]*/,
                           int *coors /*[V,3]*/, int *num_points /*[V]*/, int *p2v /*[N]*/,
                           int *p2s /*[N]*/, float *feats /*[V,F]*/, float *grad_raw /*[N,C]*/)
{
    orc_perturb_apply(points, raw, eb, fb, N, C, perturbed, NULL);
    int M = orc_hard_voxelize(perturbed, N, C, vs, pcr, P, V, table_mode, voxels, coors,
                              num_points, p2v, p2s);
    if (M < 0) return M;
    orc_vfe_mean_fwd(voxels, num_points, M, P, C, F, feats);
    /* backward: slot gradient g/n gathered at the kept points, then through clamp and scale */
    for (int i = 0; i < N; ++i) {
        int v = p2v[i];
        for (int c = 0; c < C; ++c) {
            float g = (v >= 0 && c < F) ? grad_feats[(size_t)v * F + c] / (float)num_points[v] : 0.0f;
            float p = raw[(size_t)i * C + c] * eb[c];
            int pass = (p >= -fb[c]) && (p <= fb[c]);
            grad_raw[(size_t)i * C + c] = pass ? g * eb[c] : 0.0f;
        }
    }
    return M;
}

ORC_API int orc_abi_version(void) { return 1; }
