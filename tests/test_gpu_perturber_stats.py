"""GPU parity of the perturber's statistics kernels through the C-ABI: the std-normalise + clamp
in front of the MLP (/root/reference/models/adversarial/voxel_perturber.py:157-168) and the four regularisers with
their gradient (:268-299).  Checked against the fixtures made by the reference class (tests/golden/perturber_*.npz)
and against an fp64 evaluation of the same formulas; tolerance 1e-5 relative (fp32 outputs of fp64 sums)."""
import os

import numpy as np
import pytest
import torch

from robustpointclouds_b200 import VoxelPerturber, ops

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
RTOL = 1e-5


def _losses64(p64: torch.Tensor):
    C = p64.shape[1]
    l2 = torch.norm(p64, p=2, dim=1).mean()
    it = p64[:, 3].abs().mean() if C >= 4 else p64.sum() * 0
    return l2, it, p64.mean(dim=0).abs().mean(), p64.std(dim=0).std()


def _check_losses(p_np: np.ndarray, coef=(1.0, 2.0, 3.0, 4.0)):
    p = torch.from_numpy(p_np).to(DEV)
    losses, consts = ops.perturb_losses_fwd(p)
    p64 = p.double().requires_grad_(True)
    ref = _losses64(p64)
    got = losses.cpu().numpy()
    for a, b in zip(got, ref):
        assert a == pytest.approx(float(b.detach()), rel=RTOL, abs=1e-12)
    sum(c * r for c, r in zip(coef, ref)).backward()
    g = ops.perturb_losses_bwd(p, consts, torch.tensor(coef, dtype=torch.float32, device=DEV)).cpu().numpy()
    refg = torch.nan_to_num(p64.grad, nan=0.0).cpu().numpy()        # fp64 std backward divides 0 by 0 only where
    scale = max(float(np.abs(refg).max()), 1e-30)                   # torch's fp32 rule masks the term to 0
    np.testing.assert_allclose(g, refg, rtol=1e-4, atol=RTOL * scale)
    return got


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_losses_match_reference_fixture(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"perturber_{name}.npz"))
    got = _check_losses(g["p"])
    for v, k in zip(got, ["l2_norm", "intensity_loss", "bias_loss", "imbalance_loss"]):
        assert v == pytest.approx(float(g[f"loss_{k}"]), rel=1e-5, abs=1e-9)
    # the fixture's gradient on p is w (through y = x + p) + 1, 2, 3, 4 times the regularisers'
    p = torch.from_numpy(g["p"]).to(DEV)
    losses, consts = ops.perturb_losses_fwd(p)
    gp = ops.perturb_losses_bwd(p, consts, torch.tensor([1.0, 2.0, 3.0, 4.0], device=DEV)).cpu().numpy()
    ref = g["grad_p"] - g["w"]
    np.testing.assert_allclose(gp, ref, rtol=1e-4, atol=2e-5 * float(np.abs(g["grad_p"]).max()))


@pytest.mark.parametrize("C", [4, 5, 6, 8])
@pytest.mark.parametrize("K", [2, 33, 1025, 200_003])
def test_losses_random(K, C):
    rng = np.random.default_rng(K + C)
    p = (rng.normal(size=(K, C)) * np.array([0.2, 0.25, 0.15, 0.02, 0, 0.1, 0.3, 1.0])[:C] + 0.01).astype(np.float32)
    p[rng.random(K) < 0.1] = 0.0                                     # zero rows: norm backward is 0 there
    _check_losses(p)


def test_losses_constant_columns_and_zero_imbalance():
    p = np.zeros((100, 5), np.float32)
    _check_losses(p)                                                 # everything 0, gradient 0, no NaN
    p[:, :4] = np.random.default_rng(0).normal(size=(100, 1)).astype(np.float32)   # equal stds: imbalance of 4 equal
    got = _check_losses(p[:, :4].copy())
    assert got[3] == pytest.approx(0.0, abs=1e-7)


@pytest.mark.parametrize("C", [4, 5])
@pytest.mark.parametrize("K", [2, 777, 300_001])
def test_normalize_matches_aten(K, C):
    rng = np.random.default_rng(K * 7 + C)
    x = (rng.normal(size=(K, C)) * np.array([30, 30, 2, 0.3, 0.2])[:C] + np.array([20, 0, -1, 0.3, 0.1])[:C]).astype(np.float32)
    x[0, 0] = 1e4                                                    # exercises the +10 clamp
    xt = torch.from_numpy(x).to(DEV)
    got = ops.normalize_fwd(xt).cpu().numpy()
    s64 = x.astype(np.float64).std(axis=0, ddof=1)
    scale = (s64.astype(np.float32) + np.float32(1e-6)).astype(np.float32)
    ref = np.clip(x / scale, -10.0, 10.0)
    np.testing.assert_allclose(got, ref, rtol=2e-7, atol=0)          # one ulp of the fp32 std
    aten = torch.clamp(xt / (torch.std(xt, dim=0, keepdim=True) + 1e-6), -10.0, 10.0).cpu().numpy()
    np.testing.assert_allclose(got, aten, rtol=1e-5, atol=0)
    assert K < 200 or got.max() == 10.0


def test_normalize_guards():
    x = torch.ones((50, 4), device=DEV)                              # std 0 -> scale 1e-6 -> clamp at 10
    assert (ops.normalize_fwd(x) == 10.0).all()
    x[3, 1] = float("inf")                                           # std NaN -> unit scaling (:161-163)
    out = ops.normalize_fwd(x)
    assert out[0, 0] == 1.0 and out[3, 1] == 10.0
    assert ops.normalize_fwd(torch.zeros((0, 4), device=DEV)).shape == (0, 4)


def test_module_fused_equals_aten_path(golden_dir):
    """VoxelPerturber(fused_stats=True) against the same module with the ATen formulas"""
    g = np.load(os.path.join(golden_dir, "perturber_kitti_train.npz"))
    x = torch.from_numpy(g["x"]).to(DEV)
    outs = []
    for fused in (True, False):
        torch.manual_seed(0)
        adv = VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]], fused_stats=fused).to(DEV)
        adv.train()
        y, ld = adv(x)
        (y.sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]).backward()
        outs.append((y.detach().cpu().numpy(), {k: float(v.detach()) for k, v in ld.items()},
                     [q.grad.cpu().numpy() for q in adv.parameters()]))
    np.testing.assert_allclose(outs[0][0], outs[1][0], rtol=1e-5, atol=1e-6)
    for k in outs[0][1]:
        assert outs[0][1][k] == pytest.approx(outs[1][1][k], rel=1e-4, abs=1e-8)
    for a, b in zip(outs[0][2], outs[1][2]):
        np.testing.assert_allclose(a, b, rtol=1e-3, atol=1e-5)


def test_device_nan_guards(golden_dir):
    """nan_guards="device": same results as the host checks, decided by a device flag (no sync inside forward)"""
    g = np.load(os.path.join(golden_dir, "perturber_kitti_train.npz"))
    x = torch.from_numpy(g["x"]).to(DEV)
    res = {}
    for mode in (True, "device"):
        torch.manual_seed(0)
        adv = VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]], nan_guards=mode).to(DEV)
        adv.train()
        y, ld = adv(x)
        res[mode] = (y.detach().cpu().numpy(), {k: float(v.detach()) for k, v in ld.items()}, adv)
    assert np.array_equal(res[True][0], res["device"][0]) and res[True][1] == res["device"][1]
    adv = res["device"][2]
    assert adv.l2_norms == [] and not adv.check_guards()            # metrics still on the device
    adv.flush_metrics()
    assert adv.l2_norms == pytest.approx(res[True][2].l2_norms) and len(adv.perturbation_stats) == 1
    # a NaN in the input: the reference returns x unchanged and a zero loss (:150-153)
    xb = x.clone()
    xb[5, 2] = float("nan")
    y, ld = adv(xb)
    assert torch.equal(torch.nan_to_num(y, nan=-7.0), torch.nan_to_num(xb, nan=-7.0))
    assert all(float(v.detach()) == 0.0 for v in ld.values())
    assert adv.check_guards()
    y2, ld2 = res[True][2](xb)
    assert y2 is xb and float(ld2) == 0.0
    # NaN weights: flagged on the device, reset by check_guards like :197
    with torch.no_grad():
        next(adv.model.parameters()).fill_(float("nan"))
    y, ld = adv(x)
    assert torch.equal(y, x) and adv.check_guards()
    assert not any(torch.isnan(q).any() for q in adv.parameters())
    y, ld = adv(x)
    assert not adv.check_guards() and float(ld["l2_norm"].detach()) > 0.0


def test_input_gradient_flows_with_fused_stats():
    """An input that requires grad (the post-VFE variant behind a learned encoder) keeps its path through the
    normalisation, the MLP and the attention, like the reference's clone / divide / clamp (voxel_perturber.py:157-168):
    fused_stats=True must give the same d/dx as the ATen formulas."""
    grads = []
    for fused in (True, False):
        torch.manual_seed(3)
        adv = VoxelPerturber(hidden_channels=[16, 32, 64], fused_stats=fused, nan_guards=False).build(4, DEV).train()
        g = torch.Generator(DEV).manual_seed(9)
        x = (torch.rand(257, 4, device=DEV, generator=g) * torch.tensor([70.0, 80.0, 4.0, 1.0], device=DEV)).requires_grad_(True)
        y, ld = adv(x)
        (y.square().sum() + sum(ld.values())).backward()
        grads.append(x.grad.clone())
    assert float((grads[0] - 2 * x.detach()).abs().max()) > 1e-3           # more than the identity path: the network's share
    torch.testing.assert_close(grads[0], grads[1], rtol=1e-4, atol=1e-5)
