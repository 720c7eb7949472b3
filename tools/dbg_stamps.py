import sys, numpy as np, torch, ctypes
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_path as orc
from robustpointclouds_b200 import PerturbCfg, VoxelCfg, _lib, ops, synth
name = sys.argv[1] if len(sys.argv) > 1 else 'kitti'
p = synth.PRESETS[name]; B = 6 if name == 'kitti' else 4
C, P, V = p['C'], p['max_num_points'], p['max_voxels'][0]
cfg = VoxelCfg.make(p['voxel_size'], p['point_cloud_range'], P, V, C, C)
dist = sys.argv[2] if len(sys.argv) > 2 else 'lidar'
clouds = synth.make_batch(name, B, dist=dist)
if len(sys.argv) > 3:
    for c in clouds: c[(c[:, :3] == 0).all(1), 0] = -1000.0   # move the null returns out of range
raws = [synth.make_raw(c, 99+i) for i, c in enumerate(clouds)]
n = [len(c) for c in clouds]
dev = torch.device('cuda:0')
eb, fb = orc.perturber_bounds(C, True); bounds = PerturbCfg.make(eb, fb)
pts = torch.from_numpy(np.concatenate(clouds)).to(dev); raw = torch.from_numpy(np.concatenate(raws)).to(dev)
out = ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, want_feats=True, want_maps=True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
_lib.profile_enable(True)
import os
for it in range(4):
    if not os.environ.get('NOFLUSH'): flush.zero_()
    ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, out=out)
    torch.cuda.synchronize()
NC = int(os.environ.get('NCTA', '296'))
off = _lib.lib().rpc_workspace_stamps_offset(ctypes.byref(cfg), B, _lib.counts(n))
ws = ops._default_ws._buf[dev]
raw_st = ws[off:off + 8*16*NC].view(torch.int64).cpu().numpy().reshape(NC, 16)
st = raw_st.astype(np.float64)
t0 = st[:, 0].min()
st = (st - t0) / 1e3
names = ['clear', 'insert', 'flag', 'assign', 'rank', 'gather']
for k, nm in enumerate(names):
    start = st[:, k]; arrive = st[:, 8 + k] if k < 5 else st[:, 6]
    work = arrive - start
    top = np.argsort(-work)[:8]
    print(f"{nm:8s} start {start.min():7.2f}..{start.max():7.2f}  work min {work.min():6.2f} med {np.median(work):6.2f} p90 {np.percentile(work, 90):6.2f} max {work.max():6.2f}  all-arrived {arrive.max():7.2f}  slowest " + " ".join(f"{c}:{work[c]:.1f}" for c in top))
print('insert: loads+cell of first batch done (cta-wide) med', np.median(st[:,13]), 'max', st[:,13].max())
print('end', st[:, 6].max(), 'dist', dist, 'voxels', int(out.voxel_num[-1]), 'kept', int((out.point2voxel >= 0).sum()))
