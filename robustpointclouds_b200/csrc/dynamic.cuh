// dynamic.cuh -- dynamic voxelization (per-point cell coordinates; mmcv dynamic_voxelize_forward,
// CPU rule: all three columns are -1 for a dropped point) for a batch of sweeps.
#pragma once
#include "common.cuh"
#include "hard_fwd.cuh"

namespace rpc {

// coors row = (batch, z, y, x) when NCOL == 4 (what Det3DDataPreprocessor.voxelize returns after
// padding the batch index), (z, y, x) when NCOL == 3 (what the Voxelization module returns).
// Pure streaming: one point per thread per iteration, grid-stride.
template <int C, bool PERTURB, int NCOL>
__global__ void __launch_bounds__(kThreads)
k_dynamic_voxelize(const float *__restrict__ pts, const float *__restrict__ raw,
                   const __grid_constant__ GridDesc gd, const __grid_constant__ BatchDesc bd,
                   const __grid_constant__ PerturbDesc pd, int *__restrict__ coors)
{
    const int ntotal = bd.pt_begin[bd.B];
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < ntotal; i += gridDim.x * kThreads) {
        float x, y, z;
        load_xyz<C, PERTURB>(pts, raw, pd, i, x, y, z);
        int cx = axis_cell(x, gd, 0);
        int cy = axis_cell(y, gd, 1);
        int cz = axis_cell(z, gd, 2);
        if ((cx | cy | cz) < 0) cx = cy = cz = -1;
        if constexpr (NCOL == 4) {
            int b = 0;
#pragma unroll 1
            while (b + 1 < bd.B && i >= bd.pt_begin[b + 1]) ++b;
            reinterpret_cast<int4 *>(coors)[i] = make_int4(b, cz, cy, cx);
        } else {
            int *o = coors + (int64_t)i * 3;
            o[0] = cz; o[1] = cy; o[2] = cx;
        }
    }
}

}  // namespace rpc
