"""per-CTA phase stamps of k_dyn_scatter: python tools/dbg_stamps_dyn.py [nuscenes|kitti]"""
import ctypes, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_path as orc
from robustpointclouds_b200 import PerturbCfg, VoxelCfg, _lib, ops, synth
name = sys.argv[1] if len(sys.argv) > 1 else 'nuscenes'
p = synth.PRESETS[name]; B = 6 if name == 'kitti' else 4
C = p['C']
cfg = VoxelCfg.make(p['voxel_size'], p['point_cloud_range'], 1, 1, C, C)
clouds = synth.make_batch(name, B)
raws = [synth.make_raw(c, 99 + i) for i, c in enumerate(clouds)]
n = [len(c) for c in clouds]
dev = torch.device('cuda:0')
eb, fb = orc.perturber_bounds(C, True); bounds = PerturbCfg.make(eb, fb)
pts = torch.from_numpy(np.concatenate(clouds)).to(dev); raw = torch.from_numpy(np.concatenate(raws)).to(dev)
out = ops.dynamic_scatter_mean(pts, n, cfg, raw=raw, bounds=bounds)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
_lib.profile_enable(True)
for it in range(4):
    if not os.environ.get('NOFLUSH'): flush.zero_()
    ops.dynamic_scatter_mean(pts, n, cfg, raw=raw, bounds=bounds, out=out)
    torch.cuda.synchronize()
NC = int(os.environ.get('NCTA', '296'))
off = _lib.lib().rpc_dynamic_stamps_offset(ctypes.byref(cfg), B, _lib.counts(n))
ws = ops._default_ws._buf[dev]
st = ws[off:off + 8 * 32 * NC].view(torch.int64).cpu().numpy().reshape(NC, 32).astype(np.float64)
st = (st - st[:, 0].min()) / 1e3
for k, nm in enumerate(ops.DYN_PHASES):
    start = st[:, k]; arrive = st[:, 16 + k] if k < 8 else st[:, 9]
    work = arrive - start
    print(f"{nm:12s} start {start.min():7.2f}..{start.max():7.2f}  work min {work.min():6.2f} med {np.median(work):6.2f} "
          f"max {work.max():6.2f} (cta {work.argmax()})  all-arrived {arrive.max():7.2f}")
print('end', st[:, 9].max(), 'voxels', int(out.voxel_num[-1]))
