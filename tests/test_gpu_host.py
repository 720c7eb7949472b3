"""GPU parity of the reference-facing host API: the B200 ``VoxelPerturber`` / ``perturb_voxels`` /
``HardSimpleVFE`` against fixtures produced by the REFERENCE's own ``VoxelPerturber`` class
(tests/golden/make_golden.py), and ``Voxelization`` / ``voxelize`` / ``adversarial_step`` against the CPU
oracle.  Tolerance for floats: 1e-5 relative (north_star); integers bit-exact."""
import os

import numpy as np
import pytest

from conftest import assert_pgrad_matches_fp64
import torch

import robustpointclouds_b200 as rpc
from oracle import ref_path as orc
from robustpointclouds_b200 import ops, synth

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
RTOL = 1e-5


def _close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(float(np.abs(b).max()) if b.size else 0.0, 1e-30)
    np.testing.assert_allclose(a, b, rtol=rtol, atol=rtol * scale)


def _load(golden_dir, name):
    g = dict(np.load(os.path.join(golden_dir, name)))
    state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
    return g, state


def _perturber(g, state, hidden=None):
    C = int(g["C"]) if "C" in g else g["voxels"].shape[2]
    hidden = hidden or [int(h) for h in g["hidden"]]
    adv = rpc.VoxelPerturber(hidden_channels=hidden, voxel_size=[0.05, 0.05, 0.1] if C == 4 else [0.075, 0.075, 0.2])
    adv.build(C)
    adv.load_state_dict(state)
    return adv.to(DEV)


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_voxel_perturber_matches_reference_class(golden_dir, name):
    g, state = _load(golden_dir, f"perturber_{name}.npz")
    adv = _perturber(g, state)
    adv.train(bool(g["training"]))
    x = torch.from_numpy(g["x"]).to(DEV)
    y, ld = adv(x)
    assert set(ld) == {"l2_norm", "intensity_loss", "bias_loss", "imbalance_loss"}
    _close(y.detach().cpu().numpy(), g["y"])
    for k in ld:
        _close(ld[k].item(), g[f"loss_{k}"], rtol=1e-4)            # reductions over K rows, GPU vs CPU order
    # same scalar objective as the fixture: upstream gradient w on y plus the weighted regularisers
    w = torch.from_numpy(g["w"]).to(DEV)
    total = (y * w).sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]
    total.backward()
    for k, q in adv.named_parameters():
        got = q.grad.cpu().numpy()
        assert np.abs(got).max() <= 0.1 + 1e-7                     # the +-0.1 clamp hook
        assert_pgrad_matches_fp64(got, g, k)                       # fp64 run of the reference class, 1e-4
    if bool(g["training"]):
        assert len(adv.l2_norms) == 1 and adv.l2_norms[0] == pytest.approx(float(g["loss_l2_norm"]), rel=1e-4)


@pytest.mark.parametrize("name", ["kitti_train", "nus_train"])
def test_perturb_tail_kernel_bit_exact(golden_dir, name):
    """the fused tail alone, fed the reference's own MLP output: p and y bit-exact, gradient bit-exact"""
    g, _ = _load(golden_dir, f"perturber_{name}.npz")
    C = int(g["C"])
    raw = torch.from_numpy(g["mlp_out"] * g["attn"]).to(DEV)
    x = torch.from_numpy(g["x"]).to(DEV)
    eb, fb = rpc.perturber_bounds(C, bool(g["training"]))
    cfg = rpc.PerturbCfg.make(eb.tolist(), fb.tolist())
    y, p = ops.perturb_apply_fwd(x, raw, cfg)
    assert np.array_equal(p.cpu().numpy(), g["p"]) and np.array_equal(y.cpu().numpy(), g["y"])
    gp = torch.from_numpy(g["grad_p"]).to(DEV)
    graw = ops.perturb_apply_bwd(gp, raw, cfg).cpu().numpy()
    assert np.array_equal(graw, orc.perturb_bwd(g["grad_p"], raw.cpu().numpy(), eb.numpy(), fb.numpy()))


@pytest.mark.parametrize("name,C", [("kitti", 4), ("nus", 5)])
def test_extract_feat_matches_reference_lines(golden_dir, name, C):
    """adversarial_voxelnet.py:85-117 + HardSimpleVFE, fixture made with the reference perturber"""
    g, state = _load(golden_dir, f"extract_feat_{name}.npz")
    adv = _perturber(g, state, hidden=[16, 32, 64])
    adv.train()
    voxels = torch.from_numpy(g["voxels"]).to(DEV)
    num = torch.from_numpy(g["num_points"]).to(DEV)
    vr = ops.valid_rows(voxels, num, "rowsum")
    K = vr.k()
    assert K == int(g["valid_mask"].sum())
    assert np.array_equal(vr.row_idx[:K].cpu().numpy(), np.nonzero(g["valid_mask"])[0])
    assert K < int(g["num_points"].sum())                           # the row-sum quirk drops real points
    assert ops.valid_rows(voxels, num, "count").k() == int(g["num_points"].sum())
    captured = {}

    def spy(x):
        x = x.detach().requires_grad_(True)
        captured["x"] = x
        y, ld = adv(x)
        y.retain_grad()
        captured["y"] = y
        return y, ld

    vfe = rpc.HardSimpleVFE(num_features=C)
    coors = torch.zeros((voxels.shape[0], 4), dtype=torch.int32, device=DEV)
    feats, ld = rpc.extract_voxel_features(dict(voxels=voxels, num_points=num, coors=coors), vfe, adversary=spy,
                                           training=True, epoch=3)
    perturbed, _, _ = rpc.perturb_voxels(voxels, num, lambda x: adv(x))
    _close(perturbed.detach().cpu().numpy(), g["perturbed_voxels"])
    _close(feats.detach().cpu().numpy(), g["feats"])
    for k in ld:
        _close(ld[k].item(), g[f"loss_{k}"], rtol=1e-4)
    gfe = torch.from_numpy(g["grad_feats"]).to(DEV)
    (feats * gfe).sum().backward()
    _close(captured["y"].grad.cpu().numpy(), g["grad_perturbed_valid"])
    # gate: before epoch 3 / in eval / disabled the clean voxels go straight to the encoder (:77-78)
    for kw in (dict(training=True, epoch=2), dict(training=False, epoch=5), dict(training=True, epoch=5, disabled=True)):
        f2, ld2 = rpc.extract_voxel_features(dict(voxels=voxels, num_points=num, coors=coors), vfe, adversary=spy, **kw)
        assert ld2 is None
        assert np.array_equal(f2.cpu().numpy(), orc.vfe_mean_fwd(g["voxels"], g["num_points"], C))
    # failure paths (adversarial_voxelnet.py:118-132): clean voxels + the zero-valued four-key dict that loss() indexes
    clean_feats = orc.vfe_mean_fwd(g["voxels"], g["num_points"], C)

    def check_fallback(f3, ld3):
        assert set(ld3) == {"l2_norm", "intensity_loss", "bias_loss", "imbalance_loss"}
        assert all(isinstance(v, torch.Tensor) and v.device.type == "cuda" and float(v) == 0.0 for v in ld3.values())
        assert np.array_equal(f3.cpu().numpy(), clean_feats)

    def raises(x):                                                    # any Exception, not only RuntimeError
        raise ValueError("adversary exploded")

    check_fallback(*rpc.extract_voxel_features(dict(voxels=voxels, num_points=num, coors=coors), vfe, adversary=raises,
                                               training=True, epoch=3))
    # the perturber's own NaN fall-back returns (input, tensor(0.)): the reference's loss_dict['l2_norm'] lookup raises
    check_fallback(*rpc.extract_voxel_features(dict(voxels=voxels, num_points=num, coors=coors), vfe,
                                               adversary=lambda x: (x, torch.tensor(0.0, device=x.device)),
                                               training=True, epoch=3))
    # empty valid set: every row sums to zero (:118-121)
    zero_vox = torch.zeros_like(voxels)
    f4, ld4 = rpc.extract_voxel_features(dict(voxels=zero_vox, num_points=num, coors=coors), vfe, adversary=raises,
                                         training=True, epoch=3)
    assert set(ld4) == {"l2_norm", "intensity_loss", "bias_loss", "imbalance_loss"} and all(float(v) == 0.0 for v in ld4.values())
    assert not f4.cpu().numpy().any()


def test_valid_rows_scatter_gather_vs_oracle():
    rng = np.random.default_rng(11)
    for (M, P, C) in [(1, 1, 4), (37, 5, 4), (3000, 10, 5), (20000, 5, 4)]:
        num = rng.integers(1, P + 1, M).astype(np.int32)
        vox = rng.normal(size=(M, P, C)).astype(np.float32)
        vox[np.arange(P)[None, :] >= num[:, None]] = 0
        vox[0, 0] = 0                                               # a real all-zero row
        v_t, n_t = torch.from_numpy(vox).to(DEV), torch.from_numpy(num).to(DEV)
        for mode in ("rowsum", "count"):
            mask = orc.valid_mask(vox, num, mode)
            vr = ops.valid_rows(v_t, n_t, mode)
            K = vr.k()
            assert K == int(mask.sum())
            idx = np.nonzero(mask)[0]
            assert np.array_equal(vr.row_idx[:K].cpu().numpy(), idx)
            assert np.array_equal(vr.rows[:K].cpu().numpy(), vox.reshape(-1, C)[idx])
            r2k = np.full(M * P, -1, np.int32)
            r2k[idx] = np.arange(K)
            assert np.array_equal(vr.row2k.cpu().numpy(), r2k)
            rows = torch.from_numpy(rng.normal(size=(K, C)).astype(np.float32)).to(DEV)
            out = ops.scatter_rows(v_t, vr.row2k, rows).cpu().numpy()
            ref = vox.reshape(-1, C).copy()
            ref[idx] = rows.cpu().numpy()
            assert np.array_equal(out.reshape(-1, C), ref)
            assert np.array_equal(ops.gather_rows(torch.from_numpy(out).to(DEV), vr.row_idx, K, C).cpu().numpy(), ref[idx])


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
def test_voxelize_batch_dict_vs_oracle(preset):
    cfgd = synth.PRESETS[preset]
    clouds = [synth.make_cloud(preset, n, d, 21 + i) for i, (n, d) in
              enumerate([(20000, "lidar"), (0, "lidar"), (7001, "clustered"), (3000, "uniform")])]
    layer = rpc.Voxelization(cfgd["voxel_size"], cfgd["point_cloud_range"], cfgd["max_num_points"], (1500, 4000))
    pts = [torch.from_numpy(c).to(DEV) for c in clouds]
    for training, V in ((True, 1500), (False, 4000)):
        layer.train(training)
        d = rpc.voxelize(pts, layer, "hard")
        ref = orc.voxelize_batch(clouds, cfgd["voxel_size"], cfgd["point_cloud_range"], cfgd["max_num_points"], V, "hard")
        for k in ("voxels", "coors", "num_points"):
            assert np.array_equal(d[k].cpu().numpy(), ref[k]), k
        _close(d["voxel_centers"].cpu().numpy(), ref["voxel_centers"])
        # rpc_voxel_centers: bit-exact with the ATen expression of Det3DDataPreprocessor.voxelize
        vs = torch.tensor(cfgd["voxel_size"], device=DEV)
        lo = torch.tensor(cfgd["point_cloud_range"][:3], device=DEV)
        aten = (d["coors"][:, [3, 2, 1]].float() + 0.5) * vs + lo
        assert torch.equal(d["voxel_centers"], aten)
        assert torch.equal(ops.voxel_centers(d["coors"][:, 1:].contiguous(), layer.cfg(clouds[0].shape[1])), aten)
        lazy = rpc.voxelize(pts, layer, "hard", sync=False)              # capacity-sized, counted on the device
        m = int(lazy["voxel_num"][-1])
        assert m == d["coors"].shape[0] and torch.equal(lazy["voxel_centers"][:m], aten)
    # module API on one cloud: (voxels, coors zyx, num_points); dynamic: per-point coors
    v, c, n = layer(pts[0])
    rv, rc, rn = orc.hard_voxelize(clouds[0], cfgd["voxel_size"], cfgd["point_cloud_range"], cfgd["max_num_points"], 4000)
    assert np.array_equal(v.cpu().numpy(), rv) and np.array_equal(c.cpu().numpy(), rc) and np.array_equal(n.cpu().numpy(), rn)
    dyn = rpc.Voxelization(cfgd["voxel_size"], cfgd["point_cloud_range"], -1, -1)
    assert dyn.dynamic
    assert np.array_equal(dyn(pts[2]).cpu().numpy(), orc.dynamic_voxelize(clouds[2], cfgd["voxel_size"], cfgd["point_cloud_range"]))
    dd = rpc.voxelize(pts, dyn, "dynamic")
    rd = orc.voxelize_batch(clouds, cfgd["voxel_size"], cfgd["point_cloud_range"], -1, -1, "dynamic")
    assert np.array_equal(dd["coors"].cpu().numpy(), rd["coors"]) and np.array_equal(dd["voxels"].cpu().numpy(), rd["voxels"])


@pytest.mark.parametrize("preset", ["kitti", "nuscenes"])
def test_adversarial_step_autograd_vs_oracle(preset):
    cfgd = synth.PRESETS[preset]
    C, P, V = cfgd["C"], cfgd["max_num_points"], 3000
    clouds = [synth.make_cloud(preset, 15000, "lidar", 5), synth.make_cloud(preset, 9000, "clustered", 6)]
    raws = [synth.make_raw(c, 70 + i) for i, c in enumerate(clouds)]
    eb, fb = rpc.perturber_bounds(C, True)
    cfg = rpc.VoxelCfg.make(cfgd["voxel_size"], cfgd["point_cloud_range"], P, V, C, C)
    bounds = rpc.PerturbCfg.make(eb.tolist(), fb.tolist())
    pts = torch.from_numpy(np.concatenate(clouds)).to(DEV)
    raw = torch.from_numpy(np.concatenate(raws)).to(DEV).requires_grad_(True)
    feats, voxels, coors, num, vnum = rpc.adversarial_step(pts, raw, [len(c) for c in clouds], cfg, bounds)
    gf = np.random.default_rng(3).normal(size=tuple(feats.shape)).astype(np.float32)
    feats.backward(torch.from_numpy(gf).to(DEV))
    voff, lo = 0, 0
    vn = vnum.cpu().numpy()
    for b, (c, r) in enumerate(zip(clouds, raws)):
        M, buf = orc.step_fused(c, r, eb.numpy(), fb.numpy(), cfgd["voxel_size"], cfgd["point_cloud_range"], P, V, C,
                                np.ascontiguousarray(gf[voff:voff + V]), table_mode=1)
        assert vn[b] == M
        assert np.array_equal(voxels[voff:voff + M].cpu().numpy(), buf.voxels[:M])
        assert np.array_equal(coors[voff:voff + M, 1:].cpu().numpy(), buf.coors[:M])
        _close(feats[voff:voff + M].detach().cpu().numpy(), buf.feats[:M])
        _close(raw.grad[lo:lo + len(c)].cpu().numpy(), buf.grad_raw)
        voff += M
        lo += len(c)
