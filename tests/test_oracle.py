"""CPU tests of the ORACLE itself: pinned against fixtures generated from the reference's own
VoxelPerturber (tests/golden/make_golden.py) and against hand-derived voxelization vectors
(the reference ships no tests -- SURVEY 4 -- and mmcv is absent, so nothing upstream pins them)."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_path as orc

KITTI = dict(vs=[0.05, 0.05, 0.1], pcr=[0, -40, -3, 70.4, 40, 1], P=5, V=16000)
NUS = dict(vs=[0.075, 0.075, 0.2], pcr=[-54, -54, -5, 54, 54, 3], P=10, V=120000)


def _load(golden_dir, name):
    return dict(np.load(os.path.join(golden_dir, name)))


def _state(g):
    return {k[len("state_"):]: v for k, v in g.items() if k.startswith("state_")}


# ---------------------------------------------------------------- voxelization: hand-derived KATs
def test_grid_sizes():
    assert orc.grid_size(KITTI["vs"], KITTI["pcr"]).tolist() == [1408, 1600, 40]
    assert orc.grid_size(NUS["vs"], NUS["pcr"]).tolist() == [1440, 1440, 40]
    assert orc.grid_size([0.1, 0.1, 0.2], [-51.2, -51.2, -5, 51.2, 51.2, 3]).tolist() == [1024, 1024, 40]


@pytest.mark.parametrize("table_mode", [0, 1])
def test_hard_voxelize_hand_case(golden_dir, table_mode):
    g = _load(golden_dir, "voxel_micro.npz")
    assert orc.grid_size(g["vs"], g["pcr"]).tolist() == g["grid"].tolist()
    vox, coors, num, p2v, p2s = orc.hard_voxelize(g["pts"], g["vs"], g["pcr"], int(g["P"]), int(g["V"]),
                                                  table_mode, with_maps=True)
    assert np.array_equal(coors, g["coors"])
    assert np.array_equal(num, g["num"])
    assert np.array_equal(vox, g["vox"], equal_nan=True)
    assert np.array_equal(p2v, g["p2v"]) and np.array_equal(p2s, g["p2s"])
    # no caps: 4 voxels, first-occurrence order
    vox, coors, num = orc.hard_voxelize(g["pts"], g["vs"], g["pcr"], 8, 100, table_mode)
    assert np.array_equal(coors, g["coors_big"]) and np.array_equal(num, g["num_big"])


def test_dynamic_voxelize_hand_case(golden_dir):
    g = _load(golden_dir, "voxel_micro.npz")
    assert np.array_equal(orc.dynamic_voxelize(g["pts"], g["vs"], g["pcr"]), g["dyn"])


def test_dynamic_scatter_mean_hand_case(golden_dir):
    g = _load(golden_dir, "voxel_micro.npz")
    feats, oc, mp, cnt = orc.dynamic_scatter_mean(g["pts"], g["dyn"])
    assert np.array_equal(oc, g["ds_coors"]) and np.array_equal(cnt, g["ds_count"])
    assert np.array_equal(mp, g["ds_map"])
    np.testing.assert_allclose(feats, g["ds_feats"], rtol=1e-6)
    gf = np.arange(16, dtype=np.float32).reshape(4, 4)
    gp = orc.dynamic_scatter_mean_bwd(gf, mp, cnt)
    assert np.all(gp[3] == 0) and np.all(gp[9] == 0)
    np.testing.assert_array_equal(gp[0], gf[0] / np.float32(3))


def test_edge_cases():
    vs, pcr = KITTI["vs"], KITTI["pcr"]
    empty = np.zeros((0, 4), np.float32)
    v, c, n = orc.hard_voxelize(empty, vs, pcr, 5, 100)
    assert v.shape == (0, 5, 4) and c.shape == (0, 3) and n.shape == (0,)
    assert orc.dynamic_voxelize(empty, vs, pcr).shape == (0, 3)
    out = np.full((7, 4), 1e6, np.float32)                        # all out of range
    out[3] = [np.inf, 0, 0, 0]
    out[4] = [-np.inf, 0, 0, 0]
    v, c, n = orc.hard_voxelize(out, vs, pcr, 5, 100)
    assert v.shape[0] == 0
    assert np.all(orc.dynamic_voxelize(out, vs, pcr) == -1)
    # boundaries: x == lo kept (cell 0); x == hi dropped; largest float below hi kept in last cell
    hi = np.float32(70.4)
    b = np.array([[0, -40, -3, 1], [hi, 0, 0, 1], [np.nextafter(hi, np.float32(0)), 0, 0, 1],
                  [0, 40, 0, 1], [0, 0, 1, 1]], np.float32)
    d = orc.dynamic_voxelize(b, vs, pcr)
    assert d[0].tolist() == [0, 0, 0]
    assert d[1].tolist() == [-1, -1, -1]
    assert d[2][2] == 1407
    assert d[3].tolist() == [-1, -1, -1] and d[4].tolist() == [-1, -1, -1]
    # duplicates + P overflow; null returns are IN range for both datasets
    dup = np.tile(np.array([[10.01, 0.01, -1.01, 0.5]], np.float32), (9, 1))
    v, c, n = orc.hard_voxelize(dup, vs, pcr, 5, 100)
    assert n.tolist() == [5] and v.shape == (1, 5, 4)
    z = np.zeros((3, 4), np.float32)
    v, c, n = orc.hard_voxelize(z, vs, pcr, 5, 100)
    assert n.tolist() == [3] and c.tolist() == [[30, 800, 0]]


@pytest.mark.parametrize("cfg,C,N", [(KITTI, 4, 20000), (NUS, 5, 30000)])
def test_hard_voxelize_properties(cfg, C, N):
    rng = np.random.default_rng(7)
    lo, hi = np.array(cfg["pcr"][:3]), np.array(cfg["pcr"][3:])
    pts = np.zeros((N, C), np.float32)
    cen = rng.uniform(lo, hi, size=(200, 3))
    pts[:, :3] = (cen[rng.integers(0, 200, N)] + rng.normal(0, 0.15, (N, 3))).astype(np.float32)
    pts[:, 3:] = rng.uniform(0, 1, (N, C - 3))
    V = 3000
    vox, coors, num, p2v, p2s = orc.hard_voxelize(pts, cfg["vs"], cfg["pcr"], cfg["P"], V, 1, with_maps=True)
    vox0, coors0, num0 = orc.hard_voxelize(pts, cfg["vs"], cfg["pcr"], cfg["P"], V, 0)
    assert np.array_equal(vox, vox0) and np.array_equal(coors, coors0) and np.array_equal(num, num0)
    M = coors.shape[0]
    assert M == V                                             # cap reached on this cloud
    kept = p2v >= 0
    assert num.sum() == kept.sum() and num.max() <= cfg["P"] and num.min() >= 1
    dyn = orc.dynamic_voxelize(pts, cfg["vs"], cfg["pcr"])
    assert np.array_equal(dyn[kept], coors[p2v[kept]])        # every kept point sits in its voxel
    assert np.array_equal(vox[p2v[kept], p2s[kept]], pts[kept])
    first = np.full(M, N)
    np.minimum.at(first, p2v[kept], np.nonzero(kept)[0])
    assert np.all(np.diff(first) > 0)                         # first-occurrence order
    pad = np.arange(cfg["P"])[None, :] >= num[:, None]
    assert np.all(vox[pad] == 0)
    assert len({tuple(r) for r in coors.tolist()}) == M       # voxels unique


# ---------------------------------------------------------------- VFE / mask / perturb primitives
def test_vfe_mean_matches_torch():
    rng = np.random.default_rng(3)
    M, P, C = 50, 5, 4
    num = rng.integers(1, P + 1, M).astype(np.int32)
    vox = rng.normal(size=(M, P, C)).astype(np.float32)
    vox[np.arange(P)[None, :] >= num[:, None]] = 0
    t = torch.from_numpy(vox).requires_grad_(True)
    ref = t[:, :, :C].sum(dim=1) / torch.from_numpy(num).type_as(t).view(-1, 1)
    np.testing.assert_allclose(orc.vfe_mean_fwd(vox, num), ref.detach().numpy(), rtol=1e-6, atol=1e-7)
    g = rng.normal(size=(M, C)).astype(np.float32)
    ref.backward(torch.from_numpy(g))
    np.testing.assert_array_equal(orc.vfe_mean_bwd(g, num, P, C), t.grad.numpy())


def test_valid_mask_quirk(golden_dir):
    g = _load(golden_dir, "extract_feat_kitti.npz")
    m = orc.valid_mask(g["voxels"], g["num_points"], "rowsum")
    assert np.array_equal(m, g["valid_mask"])
    mc = orc.valid_mask(g["voxels"], g["num_points"], "count")
    assert mc.sum() == g["num_points"].sum() == m.sum() + 2    # null return + zero-sum point


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_perturb_apply_matches_reference(golden_dir, name):
    """C primitives vs the reference class's own tensors (raw = mlp_out * attention)."""
    g = _load(golden_dir, f"perturber_{name}.npz")
    C, training = int(g["C"]), bool(g["training"])
    eb, fb = orc.perturber_bounds(C, training)
    raw = g["mlp_out"] * g["attn"]
    y, p = orc.perturb_apply(g["x"], raw, eb, fb)
    assert np.array_equal(p, g["p"]) and np.array_equal(y, g["y"])
    ls = orc.perturb_losses(p)
    for k in ("l2_norm", "intensity_loss", "bias_loss", "imbalance_loss"):
        assert ls[k] == pytest.approx(float(g[f"loss_{k}"]), rel=1e-5, abs=1e-9)
    if C == 5:
        assert np.all(p[:, 4] == 0)


def test_bounds_table():
    eb, fb = orc.perturber_bounds(4, True)
    np.testing.assert_allclose(eb, [0.208, 0.208, 0.208, 0.2], rtol=1e-6)
    np.testing.assert_allclose(fb, [0.216, 0.216, 0.216, 0.1], rtol=1e-6)
    eb, fb = orc.perturber_bounds(4, False)
    np.testing.assert_allclose(eb, [1.5666667] * 3 + [1.5], rtol=1e-6)
    np.testing.assert_allclose(fb, [5, 5, 5, 2], rtol=1e-6)
    eb, fb = orc.perturber_bounds(5, True)
    assert eb.tolist() == fb.tolist() and eb[4] == 0 and eb[0] == np.float32(0.2)
    with pytest.raises(UnboundLocalError):
        orc.perturber_bounds(3, True)


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_perturber_forward_restatement(golden_dir, name):
    """torch-CPU restatement of VoxelPerturber.forward vs the real class (same weights, same x)."""
    g = _load(golden_dir, f"perturber_{name}.npz")
    C, training = int(g["C"]), bool(g["training"])
    model, attn = orc.build_mlp(C, g["hidden"].tolist(), _state(g))
    model.train(training), attn.train(training)
    x = torch.from_numpy(g["x"])
    y, ld, p = orc.perturber_forward(model, attn, x, training)
    np.testing.assert_allclose(y.detach().numpy(), g["y"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(p.detach().numpy(), g["p"], rtol=1e-5, atol=1e-8)
    for k in ld:
        assert float(ld[k].detach()) == pytest.approx(float(g[f"loss_{k}"]), rel=1e-5, abs=1e-9)
    # backward through clamp/scale: C primitive vs autograd of the reference
    p.retain_grad()
    w = torch.from_numpy(g["w"])
    (y * w).sum().add(ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"]
                      + 4 * ld["imbalance_loss"]).backward()
    np.testing.assert_allclose(p.grad.numpy(), g["grad_p"], rtol=1e-4, atol=1e-7)


@pytest.mark.parametrize("name,C", [("kitti", 4), ("nus", 5)])
def test_extract_feat_restatement(golden_dir, name, C):
    g = _load(golden_dir, f"extract_feat_{name}.npz")
    model, attn = orc.build_mlp(C, [16, 32, 64], _state(g))
    model.train(), attn.train()
    captured = {}

    def adversary(v):
        y, ld, p = orc.perturber_forward(model, attn, v, True)
        y.retain_grad()
        captured["y"] = y
        return y, ld

    feats, ld, pert, mask = orc.extract_feat_ref(torch.from_numpy(g["voxels"]),
                                                 torch.from_numpy(g["num_points"]), adversary)
    assert np.array_equal(mask.numpy(), g["valid_mask"])
    np.testing.assert_allclose(pert.detach().numpy(), g["perturbed_voxels"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(feats.detach().numpy(), g["feats"], rtol=1e-6, atol=1e-7)
    (feats * torch.from_numpy(g["grad_feats"])).sum().backward()
    np.testing.assert_allclose(captured["y"].grad.numpy(), g["grad_perturbed_valid"], rtol=1e-6)
    # C oracle of the same backward: slot gradient g/n gathered at the valid rows
    gv = orc.vfe_mean_bwd(g["grad_feats"], g["num_points"], g["voxels"].shape[1], C)
    np.testing.assert_array_equal(gv.reshape(-1, C)[g["valid_mask"]], g["grad_perturbed_valid"])


def test_step_fused_composition():
    """orc_step_fused == composition of the primitives (north-star order)."""
    rng = np.random.default_rng(11)
    N, C = 5000, 4
    pts = np.zeros((N, C), np.float32)
    pts[:, :3] = rng.uniform([0, -40, -3], [70.4, 40, 1], (N, 3))
    pts[:, 3] = rng.uniform(0, 1, N)
    pts[:2000, :3] = pts[2000:4000, :3] + rng.normal(0, 0.02, (2000, 3)).astype(np.float32)
    raw = rng.uniform(-1.5, 1.5, (N, C)).astype(np.float32)
    eb, fb = orc.perturber_bounds(C, True)
    V, P = 1500, 5
    gf = rng.normal(size=(V, C)).astype(np.float32)
    M, buf = orc.step_fused(pts, raw, eb, fb, KITTI["vs"], KITTI["pcr"], P, V, C, gf, table_mode=1)
    y, p = orc.perturb_apply(pts, raw, eb, fb)
    vox, coors, num, p2v, p2s = orc.hard_voxelize(y, KITTI["vs"], KITTI["pcr"], P, V, 1, with_maps=True)
    assert M == coors.shape[0]
    assert np.array_equal(buf.voxels[:M], vox) and np.array_equal(buf.coors[:M], coors)
    assert np.array_equal(buf.feats[:M], orc.vfe_mean_fwd(vox, num))
    gslots = orc.vfe_mean_bwd(gf[:M], num, P, C)
    gp = np.zeros((N, C), np.float32)
    k = p2v >= 0
    gp[k] = gslots[p2v[k], p2s[k]]
    assert np.array_equal(buf.grad_raw, orc.perturb_bwd(gp, raw, eb, fb))


def _xn_of(x: np.ndarray) -> np.ndarray:
    import torch
    xt = torch.from_numpy(x)
    return torch.clamp(xt / (torch.std(xt, dim=0, keepdim=True) + 1e-6), -10.0, 10.0).numpy()


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_attention_gate_restatement_vs_reference_class(golden_dir, name):
    """oracle attention_gate against the gate and the MLP output recorded from the reference's own class"""
    g = np.load(os.path.join(golden_dir, f"perturber_{name}.npz"))
    xn = _xn_of(g["x"])
    w = [g[f"state_attention.{k}"] for k in ("0.weight", "0.bias", "2.weight", "2.bias")]
    out, gate = orc.attention_gate(xn, g["mlp_out"], *w)
    np.testing.assert_allclose(gate, g["attn"], rtol=1e-5, atol=1e-7)
    # p = clamp(raw * eb, +-fb): the fixture's p pins the gated product through the tail
    eb, fb = orc.perturber_bounds(int(g["C"]), bool(g["training"]))
    p = np.clip(out * eb[None, :], -fb[None, :], fb[None, :])
    np.testing.assert_allclose(p, g["p"], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("C", [1, 3, 4, 5, 8])
def test_attention_gate_bwd_restatement_vs_autograd(C):
    import torch
    rng = np.random.default_rng(C)
    K, mid = 257, max(C // 2, 1)
    xn, t, go = (rng.standard_normal((K, C)) for _ in range(3))
    w1, b1, w2, b2 = rng.standard_normal((mid, C)), rng.standard_normal(mid), rng.standard_normal((1, mid)), rng.standard_normal(1)
    tt = [torch.tensor(v, dtype=torch.float64, requires_grad=True) for v in (xn, t, w1, b1, w2, b2)]
    a = torch.sigmoid(torch.relu(tt[0] @ tt[2].T + tt[3]) @ tt[4].T + tt[5])
    (tt[1] * a * torch.tensor(go)).sum().backward()
    r = orc.attention_gate_bwd(go, xn, t, w1, b1, w2, b2)
    for key, q in zip(("grad_xn", "grad_t", "grad_w1", "grad_b1", "grad_w2", "grad_b2"), tt):
        np.testing.assert_allclose(r[key], q.grad.numpy(), rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("M,N", [(64, 4), (4, 64), (16, 32)])
def test_linear_wgrad_restatement_vs_autograd(M, N):
    rng = np.random.default_rng(M * N)
    K = 301
    x, go = rng.standard_normal((K, N)), rng.standard_normal((K, M))
    lin = torch.nn.Linear(N, M).double()
    lin(torch.tensor(x)).backward(torch.tensor(go))
    gw, gb = orc.linear_wgrad(go, x)
    np.testing.assert_allclose(gw, lin.weight.grad.numpy(), rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(gb, lin.bias.grad.numpy(), rtol=1e-10, atol=1e-12)
