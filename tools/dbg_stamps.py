"""Per-CTA phase stamps of k_hard_fwd (setup / insert / barrier / scan / emit / tail) on one workload:
python tools/dbg_stamps.py [kitti|nuscenes] [lidar|uniform|clustered]
env NOFLUSH=1: warm L2; NCTA: grid size; BATCH / POINTS: batch shape; RPC_SYNTH_NULL=0: no null returns"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from robustpointclouds_b200 import PerturbCfg, VoxelCfg, _lib, ops, synth  # noqa: E402
from robustpointclouds_b200.voxel_perturber import perturber_bounds  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "kitti"
dist = sys.argv[2] if len(sys.argv) > 2 else "lidar"
p = synth.PRESETS[name]
B = int(os.environ.get("BATCH", "6" if name == "kitti" else "4"))
N = int(os.environ.get("POINTS", str(p["N"])))
C, P, V = p["C"], p["max_num_points"], p["max_voxels"][0]
cfg = VoxelCfg.make(p["voxel_size"], p["point_cloud_range"], P, V, C, C)
clouds = synth.make_batch(name, B, N=N, dist=dist)
raws = [synth.make_raw(c, 99 + i) for i, c in enumerate(clouds)]
n = [len(c) for c in clouds]
dev = torch.device("cuda:0")
eb, fb = perturber_bounds(C, True)
bounds = PerturbCfg.make(eb.tolist(), fb.tolist())
pts = torch.from_numpy(np.concatenate(clouds)).to(dev)
raw = torch.from_numpy(np.concatenate(raws)).to(dev)
out = ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, want_feats=True, want_maps=True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
_lib.profile_enable(True)
for it in range(4):
    if not os.environ.get("NOFLUSH"):
        flush.zero_()
    ops.hard_voxelize(pts, n, cfg, raw=raw, bounds=bounds, out=out)
    torch.cuda.synchronize()
NC = int(os.environ.get("NCTA", "296"))
off = _lib.lib().rpc_workspace_stamps_offset(ctypes.byref(cfg), B, _lib.counts(n))
ws = ops._hard_ws.current(dev)
raw_st = ws[off:off + 8 * 16 * NC].view(torch.int64).cpu().numpy().reshape(NC, 16)
st = raw_st.astype(np.float64)
t0 = st[:, 0].min()
st = (st - t0) / 1e3
names = ["setup", "insert", "barrier", "scan", "emit", "tail"]
for k, nm in enumerate(names):
    start, end = st[:, k], st[:, k + 1]
    work = end - start
    top = np.argsort(-work)[:6]
    print(f"{nm:8s} start {start.min():7.2f}..{start.max():7.2f}  dur min {work.min():6.2f} med {np.median(work):6.2f} "
          f"p90 {np.percentile(work, 90):6.2f} max {work.max():6.2f}  all-done {end.max():7.2f}  slowest " +
          " ".join(f"{c}:{work[c]:.1f}" for c in top))
smid = raw_st[:, 7]
ins = st[:, 2] - st[:, 1]
order = np.argsort(ins)
print("insert by CTA (fastest 8): " + " ".join(f"{c}@sm{smid[c]}:{ins[c]:.1f}" for c in order[:8]))
print("insert by CTA (slowest 8): " + " ".join(f"{c}@sm{smid[c]}:{ins[c]:.1f}" for c in order[-8:]))
if os.environ.get("PERSM"):
    persm = {}
    for c in range(NC):
        persm.setdefault(int(smid[c]), []).append(ins[c])
    print("insert mean per SM (sm:us): " + " ".join(f"{m}:{np.mean(persm[m]):.0f}" for m in sorted(persm)))
if int(os.environ.get("RPC_DBG", "0")) & 64:
    nt = sum((x + 255) // 256 for x in n)
    rec = ws[off + 8 * 16 * 296: off + 8 * 16 * 296 + 24 * nt].view(torch.int64).cpu().numpy().reshape(nt, 3)
    t_s = (rec[:, 0] - t0) / 1e3; t_e = (rec[:, 1] - t0) / 1e3; cta = rec[:, 2] >> 32; inr = rec[:, 2] & 0xFFFFFFFF
    dur = t_e - t_s
    for lo, hi in ((0, 0), (1, 15), (16, 31), (32, 32)):
        sel = (inr >= lo) & (inr <= hi)
        if sel.any():
            print(f"tiles with {lo}-{hi} in-range points in warp 0: n={sel.sum():5d} dur med {np.median(dur[sel]):6.2f} p90 {np.percentile(dur[sel], 90):6.2f} max {dur[sel].max():6.2f}")
    order_t = np.argsort(t_s)
    for a in range(0, nt, max(nt // 12, 1)):
        sl = order_t[a:a + max(nt // 12, 1)]
        print(f"  tiles started {t_s[sl].min():6.1f}..{t_s[sl].max():6.1f} us: dur med {np.median(dur[sl]):6.2f} max {dur[sl].max():6.2f}; in-range share {np.mean(inr[sl] > 0):.2f}")
    recb = ws[off + 8 * 16 * 296 + 24 * 5000: off + 8 * 16 * 296 + 24 * (5000 + nt)].view(torch.int64).cpu().numpy().reshape(nt, 3)
    tb_s = (recb[:, 0] - t0) / 1e3; tb_e = (recb[:, 1] - t0) / 1e3; candb = recb[:, 2] & 0xFFFFFFFF
    durb = tb_e - tb_s
    print("emit tiles: candidates in warp 0 == 0: n=%d dur med %.2f max %.2f | > 0: n=%d dur med %.2f p90 %.2f max %.2f" % (
        (candb == 0).sum(), np.median(durb[candb == 0]), durb[candb == 0].max(), (candb > 0).sum(),
        np.median(durb[candb > 0]), np.percentile(durb[candb > 0], 90), durb[candb > 0].max()))
    ob = np.argsort(tb_s)
    for a in range(0, nt, max(nt // 8, 1)):
        sl = ob[a:a + max(nt // 8, 1)]
        print(f"  emit tiles started {tb_s[sl].min():6.1f}..{tb_s[sl].max():6.1f} us: dur med {np.median(durb[sl]):6.2f} max {durb[sl].max():6.2f}; tile ids {sl.min()}..{sl.max()}")
    heavy = np.argsort(-durb)[:10]
    print("  slowest emit tiles (tile:dur@start):", " ".join(f"{t}:{durb[t]:.1f}@{tb_s[t]:.0f}" for t in heavy))
    persm = {}
    for i in range(nt):
        persm.setdefault(int(smid[cta[i]]), []).append(dur[i] if inr[i] > 0 else np.nan)
    v = np.array([np.nanmedian(persm[k]) if np.isfinite(persm[k]).any() else np.nan for k in sorted(persm)])
    print("median duration of in-range tiles per SM: min", np.nanmin(v), "med", np.nanmedian(v), "max", np.nanmax(v))
print("probes counted (RPC_DBG=16), per call:", int(ws[28:32].view(torch.int32).item()) / 5.0, "points", sum(n))
if int(os.environ.get("RPC_DBG", "0")) & 65536:
    wx = ws[off + 8 * (16 * 296 + 30000): off + 8 * (16 * 296 + 30000 + 64 * 32)].view(torch.int64).cpu().numpy().reshape(64, 32)
    wx = (wx - t0) / 1e3 - st[:64, 4:5]
    print("warp exit from the emit loop, relative to the CTA's emit start (us): per-CTA min med max over warps, first 6 CTAs:")
    for c in range(6):
        print("   CTA", c, np.round(np.sort(wx[c]), 2))
rounds, rpts = raw_st[:, 8], raw_st[:, 9]
print("rounds per CTA: min %d med %d max %d; points per CTA: min %d med %d max %d; corr(insert time, rounds) %.2f, corr(insert time, points) %.2f" % (
    rounds.min(), np.median(rounds), rounds.max(), rpts.min(), np.median(rpts), rpts.max(),
    np.corrcoef(ins, rounds)[0, 1], np.corrcoef(ins, rpts)[0, 1]))
print("slowest CTAs: " + " ".join(f"{c}:{ins[c]:.1f}us/{rounds[c]}r/{rpts[c]}p" for c in order[-8:]) + " | fastest: " + " ".join(f"{c}:{ins[c]:.1f}us/{rounds[c]}r/{rpts[c]}p" for c in order[:6]))
print("CTAs that found the workspace clean:", int(raw_st[:, 15].sum()), "of", NC)
print("end", st[:, 6].max(), "dist", dist, "voxels", int(out.voxel_num[-1]), "kept", int((out.point2voxel >= 0).sum()))
