"""GPU parity of the dynamic configuration (BASELINE.json config 4): dynamic voxelization + DynamicScatter(mean)
forward and its backward to the raw perturbation, called through the C-ABI, against the CPU oracle.
Per-point coors, voxel order (sorted (z, y, x) per sample), point->voxel map and counts are bit-exact;
means and gradients within 1e-5 relative (the GPU sums with fp32 red.add, order-free; an fp64 shadow of the
same sums arbitrates)."""
import numpy as np
import pytest
import torch

from oracle import ref_path as orc
from robustpointclouds_b200 import DynamicScatter, PerturbCfg, VoxelCfg, dynamic_step, ops, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _oracle_dynamic(clouds, raws, eb, fb, vs, pcr, F, grad_feats):
    feats, vco, cnt, mp, coo, graw, f64 = [], [], [], [], [], [], []
    voff = 0
    for b, pts in enumerate(clouds):
        y = pts
        if raws is not None and len(pts):
            y, _ = orc.perturb_apply(pts, raws[b], eb, fb)
        c = orc.dynamic_voxelize(y, vs, pcr) if len(pts) else np.zeros((0, 3), np.int32)
        f, oc, m, n = orc.dynamic_scatter_mean(y, c) if len(pts) else (
            np.zeros((0, pts.shape[1]), np.float32), np.zeros((0, 3), np.int32), np.zeros((0,), np.int32),
            np.zeros((0,), np.int32))
        M = len(n)
        s64 = np.zeros((M, pts.shape[1]), np.float64)
        k = m >= 0
        np.add.at(s64, m[k], y[k].astype(np.float64))
        f64.append(s64[:, :F] / np.maximum(n, 1)[:, None])
        g = orc.dynamic_scatter_mean_bwd(np.ascontiguousarray(grad_feats[voff:voff + M]), m, n) if len(pts) else \
            np.zeros_like(pts)
        if g.shape[1] < pts.shape[1]:
            g = np.concatenate([g, np.zeros((len(pts), pts.shape[1] - g.shape[1]), np.float32)], axis=1)
        if raws is not None and len(pts):
            g = orc.perturb_bwd(g, raws[b], eb, fb)
        feats.append(f[:, :F]); cnt.append(n); graw.append(g)
        vco.append(np.concatenate([np.full((M, 1), b, np.int32), oc], axis=1))
        coo.append(np.concatenate([np.full((len(pts), 1), b, np.int32), c], axis=1))
        mp.append(np.where(k, m + voff, -1).astype(np.int32))
        voff += M
    cat = np.concatenate
    return dict(feats=cat(feats), f64=cat(f64), vcoors=cat(vco), count=cat(cnt), map=cat(mp), coors=cat(coo),
                grad_raw=cat(graw), voxel_num=[len(n) for n in cnt])


def _run(clouds, preset, perturb, F=None, seed=11):
    cfgd = synth.PRESETS[preset]
    vs, pcr, C = cfgd["voxel_size"], cfgd["point_cloud_range"], cfgd["C"]
    F = F or C
    dev = torch.device("cuda:0")
    cfg = VoxelCfg.make(vs, pcr, 1, 1, C, F)
    n = [len(c) for c in clouds]
    N = sum(n)
    pts = torch.from_numpy(np.concatenate(clouds)).to(dev)
    raws = raw_t = bounds = eb = fb = None
    if perturb:
        eb, fb = orc.perturber_bounds(C, True)
        raws = [synth.make_raw(c, seed + i) for i, c in enumerate(clouds)]
        raw_t = torch.from_numpy(np.concatenate(raws)).to(dev)
        bounds = PerturbCfg.make(eb, fb)
    out = ops.dynamic_scatter_mean(pts, n, cfg, raw=raw_t, bounds=bounds)
    gf = np.random.default_rng(seed).normal(size=(max(N, 1), F)).astype(np.float32)
    graw = ops.step_bwd(torch.from_numpy(gf).to(dev), out.num_points, out.point2voxel, cfg, raw=raw_t, bounds=bounds)
    torch.cuda.synchronize()
    ref = _oracle_dynamic(clouds, raws, eb, fb, vs, pcr, F, gf)
    vn = out.voxel_num.cpu().numpy()
    assert vn[:-1].tolist() == ref["voxel_num"] and vn[-1] == sum(ref["voxel_num"])
    M = int(vn[-1])
    assert np.array_equal(out.coors.cpu().numpy(), ref["coors"])
    assert np.array_equal(out.voxel_coors[:M].cpu().numpy(), ref["vcoors"])
    assert np.array_equal(out.num_points[:M].cpu().numpy(), ref["count"])
    assert np.array_equal(out.point2voxel.cpu().numpy(), ref["map"])
    got = out.feats[:M].cpu().numpy()
    scale = max(float(np.abs(ref["f64"]).max()) if M else 0.0, 1e-30)
    np.testing.assert_allclose(got, ref["f64"], rtol=RTOL, atol=RTOL * scale)
    np.testing.assert_allclose(got, ref["feats"], rtol=RTOL, atol=RTOL * scale)
    gs = max(float(np.abs(ref["grad_raw"]).max()) if N else 0.0, 1e-30)
    np.testing.assert_allclose(graw.cpu().numpy(), ref["grad_raw"], rtol=RTOL, atol=RTOL * gs)
    return M


@pytest.mark.parametrize("preset,N", [("kitti", 30000), ("nuscenes", 40000)])
@pytest.mark.parametrize("dist", ["lidar", "uniform", "clustered"])
@pytest.mark.parametrize("perturb", [False, True])
def test_dynamic_scatter_parity_single(preset, N, dist, perturb):
    assert _run([synth.make_cloud(preset, N, dist, 1234)], preset, perturb) > 0


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
def test_dynamic_scatter_ragged_batch(preset):
    C = synth.PRESETS[preset]["C"]
    clouds = [synth.make_cloud(preset, 20001, "lidar", 1), np.zeros((0, C), np.float32),
              synth.make_cloud(preset, 1023, "clustered", 2), np.full((77, C), 1e6, np.float32),
              synth.make_cloud(preset, 4097, "uniform", 3)]
    assert _run(clouds, preset, True) > 0


def test_dynamic_scatter_num_features_subset():
    assert _run([synth.make_cloud("nuscenes", 9000, "lidar", 5)], "nuscenes", True, F=4) > 0


def test_dynamic_scatter_full_size():
    """config 4 at full size: 4 nuScenes sweeps of 300k points"""
    assert _run(synth.make_batch("nuscenes", 4), "nuscenes", True) > 0


@pytest.mark.parametrize("dist", ["lidar", "uniform"])
def test_dynamic_scatter_sweep_top(dist):
    """top of the configs[4] sweep: 2 sweeps of 2 M points"""
    assert _run([synth.make_cloud("nuscenes", 2_000_000, dist, 60 + i) for i in range(2)], "nuscenes", True) > 0


def test_dynamic_scatter_empty_and_all_dropped():
    dev = torch.device("cuda:0")
    cfgd = synth.NUSCENES
    cfg = VoxelCfg.make(cfgd["voxel_size"], cfgd["point_cloud_range"], 1, 1, 5, 5)
    out = ops.dynamic_scatter_mean(torch.zeros((0, 5), device=dev), [0], cfg)
    assert out.voxel_num.cpu().tolist() == [0, 0]
    far = torch.full((100, 5), 1e6, device=dev)
    out = ops.dynamic_scatter_mean(far, [60, 40], cfg)
    assert out.voxel_num.cpu().tolist() == [0, 0, 0]
    assert (out.point2voxel.cpu().numpy() == -1).all()
    assert (out.coors[:, 1:].cpu().numpy() == -1).all()


def test_dynamic_scatter_module_and_autograd():
    """mmcv DynamicScatter(points, coors) surface and the fused dynamic_step autograd"""
    dev = torch.device("cuda:0")
    cfgd = synth.KITTI
    vs, pcr = cfgd["voxel_size"], cfgd["point_cloud_range"]
    clouds = [synth.make_cloud("kitti", 5000, "lidar", 3), synth.make_cloud("kitti", 3000, "clustered", 4)]
    ref_co = orc.voxelize_batch(clouds, vs, pcr, -1, -1, "dynamic")["coors"]
    pts = torch.from_numpy(np.concatenate(clouds)).to(dev).requires_grad_(True)
    scatter = DynamicScatter(vs, pcr, True)
    f, c = scatter(pts, torch.from_numpy(ref_co).to(dev))
    ref = _oracle_dynamic(clouds, None, None, None, vs, pcr, 4, np.ones((8000, 4), np.float32))
    assert np.array_equal(c.cpu().numpy(), ref["vcoors"])
    np.testing.assert_allclose(f.detach().cpu().numpy(), ref["f64"], rtol=RTOL, atol=RTOL * 80)
    f.sum().backward()
    np.testing.assert_allclose(pts.grad.cpu().numpy(), ref["grad_raw"], rtol=RTOL, atol=1e-7)
    # single-sample (z, y, x) form
    f1, c1 = scatter(pts[:5000].detach(), torch.from_numpy(ref_co[:5000, 1:]).to(dev))
    assert np.array_equal(c1.cpu().numpy(), ref["vcoors"][: c1.shape[0], 1:])
    # fused step with autograd
    eb, fb = orc.perturber_bounds(4, True)
    raws = [synth.make_raw(x, 20 + i) for i, x in enumerate(clouds)]
    raw = torch.from_numpy(np.concatenate(raws)).to(dev).requires_grad_(True)
    cfg = VoxelCfg.make(vs, pcr, 1, 1, 4, 4)
    feats, vco, num, vnum, p2v = dynamic_step(pts.detach(), raw, [5000, 3000], cfg, PerturbCfg.make(eb, fb))
    g = np.random.default_rng(1).normal(size=(8000, 4)).astype(np.float32)
    feats.backward(torch.from_numpy(g).to(dev))
    ref2 = _oracle_dynamic(clouds, raws, eb, fb, vs, pcr, 4, g)
    M = int(vnum[-1])
    assert np.array_equal(vco[:M].cpu().numpy(), ref2["vcoors"])
    np.testing.assert_allclose(raw.grad.cpu().numpy(), ref2["grad_raw"], rtol=RTOL, atol=1e-7)
