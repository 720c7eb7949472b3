// dyn_scatter.cuh -- dynamic voxelization + DynamicScatter(mean) of a batch of sweeps as ONE persistent
// cooperative kernel, sm_100a (BASELINE.json config 4; SURVEY 8a-U "DynamicScatter(mean) semantics").
//
// mmcv's DynamicScatter orders the voxels of a sample by the lexicographically sorted unique (z, y, x)
// rows (torch.unique_dim) and mmdet3d concatenates the samples in batch order.  Sorted (z, y, x) is the
// ascending linear cell id (z*gy + y)*gx + x, so no sort is needed: occupancy is recorded in a two-level
// bitmap over the global cell id  b * Gp + cell  (Gp = cells per sample rounded up to 8192) and a voxel's
// row is the number of occupied cells below it:
//
//   coarse level  one bit per GROUP of 256 consecutive cells            (B*Gp/8192 words; ~160 KB at nuScenes sizes)
//   fine level    256 bits (32 bytes, one L2 sector) per OCCUPIED group, allocated by the coarse scan
//
//   D0 clear    coarse words := 0
//   D1 mark     per point: [perturb] -> cell -> coors row out; cell id kept; coarse bit set (red.or, only
//               when a plain load shows it clear)
//   D2 coarse   chunked popcount scan of the coarse words: slot of every occupied group, group id of
//               every slot, fine groups of the chunk zeroed                               (2 sub-phases)
//   D3 fine     per point: slot of its group (coarse word + prefix), fine bit set
//   D4 rank     chunked popcount scan of the fine groups: voxel rows; per occupied cell the voxel's
//               (b, z, y, x), count := 0, feature row := 0                                (2 sub-phases)
//   D5 sum      per point: voxel row = group prefix + bits below -> point2voxel; count += 1 and feature
//               row += point (fp32 red.add: order-free, like mmcv's own CUDA reduce)
//   D6 mean     feature rows /= count
//
// The backward is k_step_bwd with point2voxel := this map and num_points := the counts.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"
#include "hard_fwd.cuh"

namespace rpc {

constexpr int kDynThreads = 1024;
constexpr int kDynCtasPerSM = 2;
constexpr int kGroupCells = 256;                  // cells per fine group (8 words, 32 bytes)
constexpr int kCoarseCells = kGroupCells * 32;    // cells per coarse word

struct DynParams {
    const float *pts, *raw;                       // [ntotal, C]
    const int *coors_in;                          // [ntotal, 4] (b, z, y, x) to scatter by, or NULL: computed
    int *coors_out;                               // [ntotal, 4] or NULL
    int *p2v;                                     // [ntotal] voxel row of each point, -1 dropped
    float *feats;                                 // [cap, F]
    int *vcoors;                                  // [cap, 4]
    int *count;                                   // [cap]
    int *voxel_num;                               // [B + 1]
    uint2 *coarse;                                // [W1] {bits, occupied groups in the words before}
    uint32_t *fine;                               // [A_cap * 8]
    uint32_t *gpre;                               // [A_cap] voxel rows before the group
    int *slot_grp;                                // [A_cap] group id of each slot
    int *cell;                                    // [ntotal] linear cell within the sample, -1 dropped
    int *pslot;                                   // [ntotal] slot of the point's group
    uint32_t *part;                               // [2 * gridDim] chunk sums of the two scans
    int *ssb;                                     // [B + 1] first slot of each sample
    long long Gp;                                 // cells per sample, padded to kCoarseCells
    int W1;                                       // coarse words
    int ntotal;
    GridDesc gd;
    BatchDesc bd;
    PerturbDesc pd;
};

// exclusive prefix of `part[0 .. nct)` at index blockIdx.x and the grand total (all CTAs compute it)
__device__ __forceinline__ uint32_t chunk_base(const uint32_t *part, int nct, int *sm, uint32_t *total)
{
    __shared__ int s_mine;
    int acc = 0;                                                   // uniform over the CTA
    for (int j0 = 0; j0 < nct; j0 += kDynThreads) {
        const int j = j0 + threadIdx.x;
        const int v = j < nct ? (int)__ldcg(part + j) : 0;
        int tot;
        const int e = block_excl_scan<kDynThreads>(v, sm, &tot);
        if (j == blockIdx.x) s_mine = acc + e;
        acc += tot;
    }
    __syncthreads();
    *total = (uint32_t)acc;
    return (uint32_t)s_mine;
}

template <int C, bool PERTURB>
__global__ void __launch_bounds__(kDynThreads, kDynCtasPerSM)
k_dyn_scatter(const __grid_constant__ DynParams p)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ int sm[33];
    const int tid = threadIdx.x, nct = gridDim.x;
    const long long gstride = (long long)nct * kDynThreads;
    const long long gtid = (long long)blockIdx.x * kDynThreads + tid;
    const int F = p.gd.F;
    const int gxy = p.gd.g[0] * p.gd.g[1];

    // ------------------------------------------------------------------------------- D0 clear
    for (long long w = gtid; w < p.W1; w += gstride) p.coarse[w] = make_uint2(0u, 0u);
    grid.sync();

    // ------------------------------------------------------------------------------- D1 mark
    for (long long i = gtid; i < p.ntotal; i += gstride) {
        int b, cell;
        if (p.coors_in) {
            const int4 c = __ldg(reinterpret_cast<const int4 *>(p.coors_in) + i);
            b = c.x;
            const bool ok = b >= 0 && b < p.bd.B && c.y >= 0 && c.y < p.gd.g[2] && c.z >= 0 && c.z < p.gd.g[1] &&
                            c.w >= 0 && c.w < p.gd.g[0];
            cell = ok ? (c.y * p.gd.g[1] + c.z) * p.gd.g[0] + c.w : -1;
            if (!ok) b = 0;
        } else {
            float x, y, z;
            load_xyz<C, PERTURB>(p.pts, p.raw, p.pd, i, x, y, z);
            const int cx = axis_cell(x, p.gd.lo[0], p.gd.vs[0], p.gd.gf[0]);
            const int cy = axis_cell(y, p.gd.lo[1], p.gd.vs[1], p.gd.gf[1]);
            const int cz = axis_cell(z, p.gd.lo[2], p.gd.vs[2], p.gd.gf[2]);
            b = 0;
#pragma unroll 1
            while (b + 1 < p.bd.B && i >= p.bd.pt_begin[b + 1]) ++b;
            const bool ok = (cx | cy | cz) >= 0;
            cell = ok ? (cz * p.gd.g[1] + cy) * p.gd.g[0] + cx : -1;
            if (p.coors_out)
                reinterpret_cast<int4 *>(p.coors_out)[i] = ok ? make_int4(b, cz, cy, cx) : make_int4(b, -1, -1, -1);
        }
        // the sample index rides in the top bits of pslot until D3 overwrites it (coors_in may carry any b)
        p.cell[i] = cell;
        p.pslot[i] = b;
        if (cell < 0) { p.p2v[i] = -1; continue; }
        const long long grp = ((long long)b * p.Gp + cell) >> 8;
        uint32_t *w = &p.coarse[grp >> 5].x;
        const uint32_t bit = 1u << (grp & 31);
        if (!(__ldcg(w) & bit)) atomicOr(w, bit);
    }
    grid.sync();

    // ------------------------------------------------------------------------------- D2 coarse scan
    const int cw = (p.W1 + nct - 1) / nct;                        // coarse words per CTA
    const int w_begin = min(blockIdx.x * cw, p.W1), w_end = min(w_begin + cw, p.W1);
    {
        int s = 0;
        for (int w = w_begin + tid; w < w_end; w += kDynThreads) s += __popc(__ldcg(&p.coarse[w].x));
        int tot;
        block_excl_scan<kDynThreads>(s, sm, &tot);
        if (tid == 0) p.part[blockIdx.x] = (uint32_t)tot;
    }
    grid.sync();
    uint32_t A;                                                    // occupied groups in the batch
    {
        uint32_t run = chunk_base(p.part, nct, sm, &A);
        const int wpb = (int)(p.Gp / kCoarseCells);                // coarse words per sample
        for (int w0 = w_begin; w0 < w_end; w0 += kDynThreads) {
            const int w = w0 + tid;
            uint32_t bits = w < w_end ? __ldcg(&p.coarse[w].x) : 0u;
            int tot;
            uint32_t e = run + (uint32_t)block_excl_scan<kDynThreads>(__popc(bits), sm, &tot);
            run += (uint32_t)tot;
            if (w < w_end) {
                p.coarse[w].y = e;
                if (w % wpb == 0) p.ssb[w / wpb] = (int)e;
                while (bits) {
                    const int k = __ffs(bits) - 1;
                    bits &= bits - 1;
                    p.slot_grp[e] = w * 32 + k;
                    uint4 *f = reinterpret_cast<uint4 *>(p.fine + (size_t)e * 8);
                    f[0] = make_uint4(0u, 0u, 0u, 0u);
                    f[1] = make_uint4(0u, 0u, 0u, 0u);
                    ++e;
                }
            }
        }
        if (blockIdx.x == 0 && tid == 0) p.ssb[p.bd.B] = (int)A;
    }
    grid.sync();

    // ------------------------------------------------------------------------------- D3 fine bits
    for (long long i = gtid; i < p.ntotal; i += gstride) {
        const int cell = __ldcg(p.cell + i);
        if (cell < 0) continue;
        const int b = __ldcg(p.pslot + i);
        const long long grp = ((long long)b * p.Gp + cell) >> 8;
        const uint2 cwv = __ldcg(&p.coarse[grp >> 5]);
        const uint32_t slot = cwv.y + __popc(cwv.x & ((1u << (grp & 31)) - 1u));
        p.pslot[i] = (int)slot;
        const int sub = cell & (kGroupCells - 1);                  // Gp is a multiple of 256
        uint32_t *w = p.fine + (size_t)slot * 8 + (sub >> 5);
        const uint32_t bit = 1u << (sub & 31);
        if (!(__ldcg(w) & bit)) atomicOr(w, bit);
    }
    grid.sync();

    // ------------------------------------------------------------------------------- D4 rank
    const uint32_t cgp = (A + nct - 1) / nct;                      // groups per CTA
    const uint32_t g_begin = min((uint32_t)blockIdx.x * cgp, A), g_end = min(g_begin + cgp, A);
    {
        int s = 0;
        for (uint32_t g = g_begin + tid; g < g_end; g += kDynThreads) {
            const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)g * 8);
            const uint4 a = __ldcg(f), c = __ldcg(f + 1);
            s += __popc(a.x) + __popc(a.y) + __popc(a.z) + __popc(a.w) + __popc(c.x) + __popc(c.y) + __popc(c.z) + __popc(c.w);
        }
        int tot;
        block_excl_scan<kDynThreads>(s, sm, &tot);
        if (tid == 0) p.part[nct + blockIdx.x] = (uint32_t)tot;
    }
    grid.sync();
    uint32_t M;
    {
        uint32_t run = chunk_base(p.part + nct, nct, sm, &M);
        for (uint32_t g0 = g_begin; g0 < g_end; g0 += kDynThreads) {
            const uint32_t g = g0 + tid;
            uint32_t wv[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
            int n = 0;
            if (g < g_end) {
                const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)g * 8);
                const uint4 a = __ldcg(f), c = __ldcg(f + 1);
                wv[0] = a.x; wv[1] = a.y; wv[2] = a.z; wv[3] = a.w; wv[4] = c.x; wv[5] = c.y; wv[6] = c.z; wv[7] = c.w;
#pragma unroll
                for (int q = 0; q < 8; ++q) n += __popc(wv[q]);
            }
            int tot;
            uint32_t e = run + (uint32_t)block_excl_scan<kDynThreads>(n, sm, &tot);
            run += (uint32_t)tot;
            if (g < g_end) {
                p.gpre[g] = e;
                const long long gc0 = (long long)__ldcg(p.slot_grp + g) * kGroupCells;
                const int b = (int)(gc0 / p.Gp);
                const int c0 = (int)(gc0 - (long long)b * p.Gp);
#pragma unroll 1
                for (int q = 0; q < 8; ++q) {
                    uint32_t bits = wv[q];
                    while (bits) {
                        const int k = __ffs(bits) - 1;
                        bits &= bits - 1;
                        const int cell = c0 + q * 32 + k;
                        const int z = cell / gxy, r = cell - z * gxy;
                        const int y = r / p.gd.g[0], x = r - y * p.gd.g[0];
                        reinterpret_cast<int4 *>(p.vcoors)[e] = make_int4(b, z, y, x);
                        p.count[e] = 0;
                        for (int c = 0; c < F; ++c) p.feats[(size_t)e * F + c] = 0.0f;
                        ++e;
                    }
                }
            }
        }
    }
    grid.sync();

    // ------------------------------------------------------------------------------- D5 sum
    if (blockIdx.x == 0 && tid <= p.bd.B) {
        // voxels before sample b = voxel rows before its first slot
        auto start = [&](int b) -> int {
            const int s = __ldcg(p.ssb + b);
            return (uint32_t)s < A ? (int)__ldcg(p.gpre + s) : (int)M;
        };
        p.voxel_num[tid] = tid < p.bd.B ? start(tid + 1) - start(tid) : (int)M;
    }
    for (long long i = gtid; i < p.ntotal; i += gstride) {
        const int cell = __ldcg(p.cell + i);
        if (cell < 0) continue;
        const uint32_t slot = (uint32_t)__ldcg(p.pslot + i);
        const int sub = cell & (kGroupCells - 1);
        const uint4 *f = reinterpret_cast<const uint4 *>(p.fine + (size_t)slot * 8);
        const uint4 a = __ldcg(f), c4 = __ldcg(f + 1);
        const uint32_t wv[8] = {a.x, a.y, a.z, a.w, c4.x, c4.y, c4.z, c4.w};
        uint32_t m = __ldcg(p.gpre + slot);
        const int wq = sub >> 5;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            if (q < wq) m += __popc(wv[q]);
            else if (q == wq) m += __popc(wv[q] & ((1u << (sub & 31)) - 1u));
        }
        p.p2v[i] = (int)m;
        atomicAdd(p.count + m, 1);
        if constexpr (C == 4) {
            float4 x = __ldg(reinterpret_cast<const float4 *>(p.pts) + i);
            if constexpr (PERTURB) {
                const float4 d = __ldg(reinterpret_cast<const float4 *>(p.raw) + i);
                x.x = __fadd_rn(x.x, perturb_delta(d.x, p.pd.eb[0], p.pd.fb[0]));
                x.y = __fadd_rn(x.y, perturb_delta(d.y, p.pd.eb[1], p.pd.fb[1]));
                x.z = __fadd_rn(x.z, perturb_delta(d.z, p.pd.eb[2], p.pd.fb[2]));
                x.w = __fadd_rn(x.w, perturb_delta(d.w, p.pd.eb[3], p.pd.fb[3]));
            }
            if (F == 4) {
                atomicAdd(reinterpret_cast<float4 *>(p.feats) + m, x);       // one 16-byte red.add
            } else {
                const float xv[4] = {x.x, x.y, x.z, x.w};
                for (int c = 0; c < F; ++c) atomicAdd(p.feats + (size_t)m * F + c, xv[c]);
            }
        } else {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                if (c >= F) break;
                float x = __ldg(p.pts + i * C + c);
                if constexpr (PERTURB) x = __fadd_rn(x, perturb_delta(__ldg(p.raw + i * C + c), p.pd.eb[c], p.pd.fb[c]));
                atomicAdd(p.feats + (size_t)m * F + c, x);
            }
        }
    }
    grid.sync();

    // ------------------------------------------------------------------------------- D6 mean
    for (long long e = gtid; e < (long long)M * F; e += gstride) {
        const long long m = e / F;
        p.feats[e] = __fdiv_rn(__ldcg(p.feats + e), (float)__ldcg(p.count + m));
    }
}

}  // namespace rpc
