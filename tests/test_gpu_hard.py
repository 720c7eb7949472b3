"""GPU parity: the sm_100a hard-voxelization / VFE / backward kernels, called through the C-ABI,
against the CPU oracle on the same seeded inputs.  Integer outputs (coors, ordering, assignment,
counts) bit-exact; VFE features and gradients within 1e-5 relative (north_star) -- in practice the
kernels use the oracle's operation order and the tests also record exact equality."""
import numpy as np
import pytest
import torch

from oracle import ref_path as orc
from robustpointclouds_b200 import PerturbCfg, VoxelCfg, ops, synth

pytestmark = pytest.mark.gpu

RTOL = 1e-5


def _close(a, b):
    scale = max(float(np.abs(b).max()) if b.size else 0.0, 1e-30)
    np.testing.assert_allclose(a, b, rtol=RTOL, atol=RTOL * scale)


def _oracle_batch(clouds, raws, eb, fb, vs, pcr, P, V, F, grad_feats):
    """per-sweep oracle, concatenated the way Det3DDataPreprocessor.voxelize concatenates"""
    vox, coo, num, p2v, p2s, feats, graw = [], [], [], [], [], [], []
    voff = 0
    for b, pts in enumerate(clouds):
        y = pts
        if raws is not None:
            y, _ = orc.perturb_apply(pts, raws[b], eb, fb) if len(pts) else (pts, None)
        v, c, n, pv, ps = orc.hard_voxelize(y, vs, pcr, P, V, 1, with_maps=True)
        M = len(n)
        f = orc.vfe_mean_fwd(v, n, F) if M else np.zeros((0, F), np.float32)
        gslots = orc.vfe_mean_bwd(grad_feats[voff:voff + M], n, P, pts.shape[1]) if M else None
        gp = np.zeros_like(pts)
        k = pv >= 0
        if M:
            gp[k] = gslots[pv[k], ps[k]]
        if raws is not None and len(pts):
            gp = orc.perturb_bwd(gp, raws[b], eb, fb)
        vox.append(v); num.append(n); feats.append(f); graw.append(gp)
        coo.append(np.concatenate([np.full((M, 1), b, np.int32), c], axis=1))
        p2v.append(np.where(k, pv + voff, -1).astype(np.int32)); p2s.append(ps)
        voff += M
    cat = np.concatenate
    return dict(voxels=cat(vox), coors=cat(coo), num_points=cat(num), p2v=cat(p2v), p2s=cat(p2s),
                feats=cat(feats), grad_raw=cat(graw), voxel_num=[len(n) for n in num])


def _run_case(clouds, preset, V, perturb, seed=5):
    cfgd = synth.PRESETS[preset]
    vs, pcr, P, C = cfgd["voxel_size"], cfgd["point_cloud_range"], cfgd["max_num_points"], cfgd["C"]
    F = C
    dev = torch.device("cuda:0")
    cfg = VoxelCfg.make(vs, pcr, P, V, C, F)
    n = [len(c) for c in clouds]
    pts = torch.from_numpy(np.concatenate(clouds)).to(dev)
    raws = bounds = raw_t = None
    eb = fb = None
    if perturb:
        eb, fb = orc.perturber_bounds(C, True)
        raws = [synth.make_raw(c, seed + i) for i, c in enumerate(clouds)]
        raw_t = torch.from_numpy(np.concatenate(raws)).to(dev)
        bounds = PerturbCfg.make(eb, fb)
    out = ops.hard_voxelize(pts, n, cfg, raw=raw_t, bounds=bounds, want_feats=True, want_maps=True)
    cap = out.capacity
    rng = np.random.default_rng(seed)
    gf = rng.normal(size=(cap, F)).astype(np.float32)
    graw = ops.step_bwd(torch.from_numpy(gf).to(dev), out.num_points, out.point2voxel, cfg, raw=raw_t,
                        bounds=bounds)
    torch.cuda.synchronize()
    ref = _oracle_batch(clouds, raws, eb, fb, vs, pcr, P, V, F, gf)
    vn = out.voxel_num.cpu().numpy()
    assert vn[:-1].tolist() == ref["voxel_num"] and vn[-1] == sum(ref["voxel_num"])
    M = int(vn[-1])
    assert np.array_equal(out.coors[:M].cpu().numpy(), ref["coors"])
    assert np.array_equal(out.num_points[:M].cpu().numpy(), ref["num_points"])
    assert np.array_equal(out.point2voxel.cpu().numpy(), ref["p2v"])
    assert np.array_equal(out.point2slot.cpu().numpy(), ref["p2s"])
    assert np.array_equal(out.voxels[:M].cpu().numpy(), ref["voxels"], equal_nan=True)
    _close(out.feats[:M].cpu().numpy(), ref["feats"])
    _close(graw.cpu().numpy(), ref["grad_raw"])
    exact = (np.array_equal(out.feats[:M].cpu().numpy(), ref["feats"])
             and np.array_equal(graw.cpu().numpy(), ref["grad_raw"]))
    return M, exact


@pytest.mark.parametrize("preset,N,V", [("kitti", 30000, 4000), ("nuscenes", 40000, 120000)])
@pytest.mark.parametrize("dist", ["lidar", "uniform", "clustered"])
@pytest.mark.parametrize("perturb", [False, True])
def test_hard_voxelize_parity_single(preset, N, V, dist, perturb):
    M, exact = _run_case([synth.make_cloud(preset, N, dist, 1234)], preset, V, perturb)
    assert M > 0 and exact


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
def test_hard_voxelize_parity_ragged_batch(preset):
    """ragged batch with an empty sweep and a sweep that is entirely out of range"""
    C = synth.PRESETS[preset]["C"]
    clouds = [synth.make_cloud(preset, 20001, "lidar", 1), np.zeros((0, C), np.float32),
              synth.make_cloud(preset, 1023, "clustered", 2), np.full((77, C), 1e6, np.float32),
              synth.make_cloud(preset, 4097, "uniform", 3)]
    M, exact = _run_case(clouds, preset, 2500, True)
    assert M > 0 and exact


@pytest.mark.parametrize("preset,B", [("kitti", 6), ("nuscenes", 4)])
def test_hard_voxelize_parity_full_size(preset, B):
    """BASELINE.json configs 2 and 3 at full size (oracle in sparse-table mode: seconds)"""
    V = synth.PRESETS[preset]["max_voxels"][0]
    M, exact = _run_case(synth.make_batch(preset, B), preset, V, True)
    assert M > 0 and exact


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
@pytest.mark.parametrize("dist", ["lidar", "uniform"])
def test_hard_voxelize_parity_sweep_top(preset, dist):
    """top of the BASELINE.json configs[4] sweep: 2 sweeps of 2 M points, every output against the oracle"""
    V = synth.PRESETS[preset]["max_voxels"][0]
    clouds = [synth.make_cloud(preset, 2_000_000, dist, 40 + i) for i in range(2)]
    M, exact = _run_case(clouds, preset, V, True)
    assert M > 0 and exact


def test_edge_semantics():
    """boundaries, NaN / inf, duplicates beyond P, null returns, cap with late points joining"""
    vs, pcr = [0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1]
    hi = np.float32(70.4)
    pts = np.array([[0, -40, -3, 1], [hi, 0, 0, 1], [np.nextafter(hi, np.float32(0)), 0, 0, 1],
                    [np.nan, 0, 0, 1], [np.inf, 0, 0, 1], [0, 0, 0, 0], [0, 0, 0, 0],
                    [10.01, 0.01, -1.01, .5]] + [[10.01, 0.01, -1.01, .25]] * 7 +
                   [[20, 1, 0, 1], [30, 1, 0, 1], [0, 0, 0, 0], [10.02, 0.02, -1.02, .9]], np.float32)
    dev = torch.device("cuda:0")
    for V in (100, 3, 1):
        for P in (5, 1, 3):
            cfg = VoxelCfg.make(vs, pcr, P, V, 4)
            out = ops.hard_voxelize(torch.from_numpy(pts).to(dev), [len(pts)], cfg, want_feats=True,
                                    want_maps=True)
            v, c, n, pv, ps = orc.hard_voxelize(pts, vs, pcr, P, V, 1, with_maps=True)
            M = out.total()
            assert M == len(n)
            assert np.array_equal(out.coors[:M, 1:].cpu().numpy(), c)
            assert np.array_equal(out.num_points[:M].cpu().numpy(), n)
            assert np.array_equal(out.voxels[:M].cpu().numpy(), v, equal_nan=True)
            assert np.array_equal(out.point2voxel.cpu().numpy(), pv)
            assert np.array_equal(out.point2slot.cpu().numpy(), ps)
    d = ops.dynamic_voxelize(torch.from_numpy(pts).to(dev), [len(pts)], VoxelCfg.make(vs, pcr, 1, 1, 4), ncol=3)
    assert np.array_equal(d.cpu().numpy(), orc.dynamic_voxelize(pts, vs, pcr))


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
def test_dynamic_voxelize_parity(preset):
    cfgd = synth.PRESETS[preset]
    clouds = [synth.make_cloud(preset, 50000, d, 7 + i) for i, d in enumerate(["lidar", "uniform", "clustered"])]
    cfg = VoxelCfg.make(cfgd["voxel_size"], cfgd["point_cloud_range"], 1, 1, cfgd["C"])
    dev = torch.device("cuda:0")
    got = ops.dynamic_voxelize(torch.from_numpy(np.concatenate(clouds)).to(dev), [len(c) for c in clouds], cfg)
    ref = orc.voxelize_batch(clouds, cfgd["voxel_size"], cfgd["point_cloud_range"], -1, -1, "dynamic")
    assert np.array_equal(got.cpu().numpy(), ref["coors"])


def test_vfe_mean_standalone():
    rng = np.random.default_rng(3)
    dev = torch.device("cuda:0")
    for (M, P, C, F) in [(1000, 5, 4, 4), (777, 10, 5, 5), (130, 10, 5, 4), (3, 32, 8, 3)]:
        num = rng.integers(1, P + 1, M).astype(np.int32)
        vox = rng.normal(size=(M, P, C)).astype(np.float32)
        vox[np.arange(P)[None, :] >= num[:, None]] = 0
        got = ops.vfe_mean_fwd(torch.from_numpy(vox).to(dev), torch.from_numpy(num).to(dev), F)
        assert np.array_equal(got.cpu().numpy(), orc.vfe_mean_fwd(vox, num, F))
        g = rng.normal(size=(M, F)).astype(np.float32)
        gb = ops.vfe_mean_bwd(torch.from_numpy(g).to(dev), torch.from_numpy(num).to(dev), P, C)
        assert np.array_equal(gb.cpu().numpy(), orc.vfe_mean_bwd(g, num, P, C))


@pytest.mark.parametrize("C,P,F,V", [(3, 4, 3, 500), (8, 32, 3, 1500), (6, 64, 6, 300), (4, 1, 4, 100000), (5, 3, 4, 7)])
def test_hard_voxelize_generic_shapes(C, P, F, V):
    """feature widths 3..8, large and unit max_num_points, num_features < C, a cap of a few voxels"""
    rng = np.random.default_rng(C * 100 + P)
    vs, pcr = [0.4, 0.4, 0.5], [-20.0, -20.0, -2.0, 20.0, 20.0, 2.0]
    clouds = []
    for n in (6000, 1, 2049):
        pts = rng.normal(0, 9, size=(n, C)).astype(np.float32)
        pts[:, 2] = rng.uniform(-3, 3, n)
        clouds.append(pts)
    dev = torch.device("cuda:0")
    cfg = VoxelCfg.make(vs, pcr, P, V, C, F)
    out = ops.hard_voxelize(torch.from_numpy(np.concatenate(clouds)).to(dev), [len(c) for c in clouds], cfg,
                            want_feats=True, want_maps=True)
    gf = rng.normal(size=(out.capacity, F)).astype(np.float32)
    g = ops.step_bwd(torch.from_numpy(gf).to(dev), out.num_points, out.point2voxel, cfg)
    ref = _oracle_batch(clouds, None, None, None, vs, pcr, P, V, F, gf)
    M = out.total()
    assert out.voxel_num.cpu().tolist()[:-1] == ref["voxel_num"] and M == sum(ref["voxel_num"])
    assert np.array_equal(out.coors[:M].cpu().numpy(), ref["coors"])
    assert np.array_equal(out.num_points[:M].cpu().numpy(), ref["num_points"])
    assert np.array_equal(out.voxels[:M].cpu().numpy(), ref["voxels"])
    assert np.array_equal(out.point2voxel.cpu().numpy(), ref["p2v"])
    assert np.array_equal(out.point2slot.cpu().numpy(), ref["p2s"])
    assert np.array_equal(out.feats[:M].cpu().numpy(), ref["feats"][:, :F])
    gref = ref["grad_raw"].copy()
    _close(g.cpu().numpy()[:, :F], gref[:, :F])
    assert not g.cpu().numpy()[:, F:].any()


def test_hard_voxelize_max_batch():
    """RPC_MAX_BATCH = 32 sweeps of ragged sizes in one launch"""
    clouds = [synth.make_cloud("kitti", 300 + 37 * i, "lidar" if i % 2 else "clustered", 900 + i) for i in range(32)]
    M, exact = _run_case(clouds, "kitti", 150, True)
    assert M > 0 and exact


def test_hard_voxelize_large_batch_matches_per_sweep_calls():
    """6 x 2M points (46 878 tiles > kMaxTiles: the assign phase scans GROUPS of tiles): the batched call returns,
    sweep by sweep, exactly what six single-sweep calls return (those take the ungrouped scan the oracle tests pin)"""
    cfgd = synth.PRESETS["kitti"]
    C, P, V = cfgd["C"], cfgd["max_num_points"], 40000
    dev = torch.device("cuda:0")
    cfg = VoxelCfg.make(cfgd["voxel_size"], cfgd["point_cloud_range"], P, V, C, C)
    N = 2_000_000
    g = torch.Generator(device="cpu").manual_seed(11)
    lo = torch.tensor(cfgd["point_cloud_range"][:3]) - 2.0
    hi = torch.tensor(cfgd["point_cloud_range"][3:]) + 2.0
    sweeps = []
    for b in range(6):
        x = torch.rand(N, C, generator=g)
        x[:, :3] = lo + x[:, :3] * (hi - lo)
        x[: N // 2, :3] = x[N // 2:, :3] + 0.01 * torch.rand(N - N // 2, 3, generator=g)   # crowded cells too
        sweeps.append(x)
    pts = torch.cat(sweeps).to(dev)
    out = ops.hard_voxelize(pts, [N] * 6, cfg, want_feats=True, want_maps=True)
    vn = out.voxel_num.cpu().tolist()
    assert vn[-1] == sum(vn[:-1]) and all(v == V for v in vn[:-1])
    row = 0
    for b in range(6):
        one = ops.hard_voxelize(pts[b * N:(b + 1) * N], [N], cfg, want_feats=True, want_maps=True)
        m = int(one.voxel_num[0])
        assert m == vn[b]
        assert torch.equal(one.voxels[:m], out.voxels[row:row + m])
        assert torch.equal(one.coors[:m, 1:], out.coors[row:row + m, 1:]) and bool((out.coors[row:row + m, 0] == b).all())
        assert torch.equal(one.num_points[:m], out.num_points[row:row + m])
        assert torch.equal(one.feats[:m], out.feats[row:row + m])
        p2v = out.point2voxel[b * N:(b + 1) * N]
        assert torch.equal(torch.where(one.point2voxel >= 0, one.point2voxel + row, one.point2voxel), p2v)
        assert torch.equal(one.point2slot, out.point2slot[b * N:(b + 1) * N])
        row += m
