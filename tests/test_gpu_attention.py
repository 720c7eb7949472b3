"""GPU parity of the fused attention gate (first slice of SURVEY 8f row 2) through the C-ABI:
``raw_perturbations * self.attention(xn)`` of /root/reference/models/adversarial/voxel_perturber.py:104-110, 203-205.
Checked against the tensors recorded from the reference's own class (``mlp_out``, ``attn``, ``pgrad_attention.*`` in
tests/golden/perturber_*.npz) and against the fp64 oracle restatement; tolerance 1e-5 relative."""
import os

import numpy as np
import pytest

from conftest import assert_pgrad_matches_fp64
import torch

from oracle import ref_path as orc
from robustpointclouds_b200 import VoxelPerturber, ops

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")
RTOL = 1e-5


def _close(a, b, rtol=RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(float(np.abs(b).max()) if b.size else 0.0, 1e-30)
    np.testing.assert_allclose(a, b.reshape(a.shape), rtol=rtol, atol=rtol * scale)


def _dev(*arrs):
    return [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(DEV) for a in arrs]


@pytest.mark.parametrize("name", ["kitti_train", "kitti_eval", "nus_train"])
def test_gate_matches_reference_class_fixture(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"perturber_{name}.npz"))
    x = torch.from_numpy(g["x"]).to(DEV)
    xn = ops.normalize_fwd(x)
    w = _dev(*[g[f"state_attention.{k}"] for k in ("0.weight", "0.bias", "2.weight", "2.bias")])
    (t,) = _dev(g["mlp_out"])
    out, gate = ops.attention_gate_fwd(xn, t, *w, want_gate=True)
    _close(gate.cpu().numpy(), g["attn"])
    _close(out.cpu().numpy(), g["mlp_out"] * g["attn"])


@pytest.mark.parametrize("K,C", [(1, 4), (31, 1), (257, 2), (1000, 3), (4099, 4), (70001, 5), (513, 6), (300, 7), (2049, 8)])
def test_gate_fwd_bwd_vs_oracle(K, C):
    rng = np.random.default_rng(100 * C + K % 97)
    mid = max(C // 2, 1)
    xn, t, go = (rng.standard_normal((K, C)).astype(np.float32) for _ in range(3))
    xn[0] = 0.0                                                  # a row that sits on the ReLU kink when b1 is 0
    w1, b1 = rng.standard_normal((mid, C)).astype(np.float32), rng.standard_normal(mid).astype(np.float32)
    w2, b2 = rng.standard_normal((1, mid)).astype(np.float32), rng.standard_normal(1).astype(np.float32)
    b1[0] = 0.0
    d = _dev(xn, t, w1, b1, w2, b2)
    out, gate = ops.attention_gate_fwd(*d, want_gate=True)
    ro, rg = orc.attention_gate(xn, t, w1, b1, w2, b2)
    _close(gate.cpu().numpy(), rg)
    _close(out.cpu().numpy(), ro)
    (dgo,) = _dev(go)
    grad_t, grad_xn, gw1, gb1, gw2, gb2 = ops.attention_gate_bwd(dgo, *d, want_grad_xn=True)
    r = orc.attention_gate_bwd(go, xn, t, w1, b1, w2, b2)
    _close(grad_t.cpu().numpy(), r["grad_t"])
    _close(grad_xn.cpu().numpy(), r["grad_xn"])
    # sums over K rows of fp32 products: the error of the per-thread fp32 partial sums, relative to the sum of magnitudes
    for got, key in ((gw1, "grad_w1"), (gb1, "grad_b1"), (gw2, "grad_w2"), (gb2, "grad_b2")):
        _close(got.cpu().numpy(), r[key], rtol=1e-4)
    assert ops.attention_gate_bwd(dgo, *d)[1] is None             # grad_xn only on request


def test_gate_empty_and_errors():
    C, mid = 4, 2
    z = torch.zeros((0, C), device=DEV)
    w = [torch.zeros((mid, C), device=DEV), torch.zeros(mid, device=DEV), torch.zeros((1, mid), device=DEV), torch.zeros(1, device=DEV)]
    assert ops.attention_gate_fwd(z, z, *w).shape == (0, C)
    gt, gx, gw1, gb1, gw2, gb2 = ops.attention_gate_bwd(z, z, z, *w)
    assert gt.shape == (0, C) and float(gw1.abs().sum() + gb1.abs().sum() + gw2.abs().sum() + gb2.abs().sum()) == 0.0
    with pytest.raises(ValueError):
        ops.attention_gate_fwd(torch.zeros((3, C), device=DEV), torch.zeros((3, C), device=DEV), w[0][:1], *w[1:])
    with pytest.raises(RuntimeError):
        ops.attention_gate_fwd(torch.zeros((3, C)), torch.zeros((3, C)), *[q.cpu() for q in w])


def test_module_fused_attention_equals_aten_modules(golden_dir):
    """VoxelPerturber(fused_attention=True) against the same module running nn.Sequential attention"""
    g = np.load(os.path.join(golden_dir, "perturber_kitti_train.npz"))
    state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
    x = torch.from_numpy(g["x"]).to(DEV)
    w = torch.from_numpy(g["w"]).to(DEV)
    res = []
    for fused in (True, False):
        adv = VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]], fused_attention=fused).build(4)
        adv.load_state_dict(state)
        adv = adv.to(DEV).train()
        y, ld = adv(x)
        ((y * w).sum() + ld["l2_norm"] + 3 * ld["bias_loss"]).backward()
        res.append((y.detach().cpu().numpy(), {k: q.grad.cpu().numpy() for k, q in adv.named_parameters()}))
    _close(res[0][0], res[1][0])
    for k in res[0][1]:
        np.testing.assert_allclose(res[0][1][k], res[1][1][k], rtol=1e-3, atol=2e-5, err_msg=k)
    # and the attention gradients against the reference class's own (both clamped to +-0.1 by the hook)
    for k in ("attention.0.weight", "attention.0.bias", "attention.2.weight", "attention.2.bias"):
        assert np.abs(res[0][1][k]).max() <= 0.1 + 1e-7


# ------------------------------------------------------------------ rpc_linear_wgrad (second slice of SURVEY 8f row 2)
@pytest.mark.parametrize("K,M,N", [(1, 64, 4), (15, 4, 64), (16, 64, 128), (4099, 128, 64), (70001, 64, 5), (70001, 5, 64),
                                   (33, 8, 16), (1000, 32, 16), (131687, 128, 128), (257, 100, 77), (300000, 64, 64)])
def test_linear_wgrad_vs_fp64(K, M, N):
    gen = torch.Generator(DEV).manual_seed(K + M + N)
    g = torch.randn((K, M), device=DEV, generator=gen)
    a = torch.randn((K, N), device=DEV, generator=gen) + 0.25
    gw, gb = ops.linear_wgrad(g, a)
    rw, rb = g.double().T @ a.double(), g.double().sum(0)
    # error of an fp32 sum of K products relative to the sum of magnitudes (what any split-K order is bounded by)
    sw = float((g.double().abs().T @ a.double().abs()).max())
    assert float((gw.double() - rw).abs().max()) <= 1e-5 * sw
    assert float((gb.double() - rb).abs().max()) <= 1e-5 * float(g.double().abs().sum(0).max())
    assert ops.linear_wgrad(g, a, want_bias=False)[1] is None


@pytest.mark.parametrize("K,M,N", [(777, 64, 4), (777, 5, 64), (2050, 64, 128)])
def test_linear_wgrad_vs_oracle(K, M, N):
    rng = np.random.default_rng(K + M)
    g, a = rng.standard_normal((K, M)).astype(np.float32), rng.standard_normal((K, N)).astype(np.float32)
    gw, gb = ops.linear_wgrad(*_dev(g, a))
    rw, rb = orc.linear_wgrad(g, a)
    _close(gw.cpu().numpy(), rw)
    _close(gb.cpu().numpy(), rb)


def test_linear_wgrad_empty_and_errors():
    gw, gb = ops.linear_wgrad(torch.zeros((0, 64), device=DEV), torch.zeros((0, 4), device=DEV))
    assert gw.shape == (64, 4) and float(gw.abs().sum()) == 0.0 and float(gb.abs().sum()) == 0.0
    with pytest.raises(Exception):
        ops.linear_wgrad(torch.zeros((8, 129), device=DEV), torch.zeros((8, 4), device=DEV))
    with pytest.raises(ValueError):
        ops.linear_wgrad(torch.zeros((8, 64), device=DEV), torch.zeros((9, 4), device=DEV))


def test_module_fused_wgrad_equals_nn_linear(golden_dir):
    """VoxelPerturber(fused_wgrad=True) against nn.Linear's own backward, and state-dict compatibility"""
    g = np.load(os.path.join(golden_dir, "perturber_kitti_train.npz"))
    state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
    x = torch.from_numpy(g["x"]).to(DEV)
    w = torch.from_numpy(g["w"]).to(DEV)
    grads = []
    for fused in (True, False):
        adv = VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]], fused_wgrad=fused, fused_trunk=False).build(4)
        assert set(adv.state_dict()) == set(state)
        adv.load_state_dict(state)
        adv = adv.to(DEV).train()
        y, ld = adv(x)
        ((y * w).sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]).backward()
        grads.append({k: q.grad.cpu().numpy() for k, q in adv.named_parameters()})
    for k in grads[0]:
        np.testing.assert_allclose(grads[0][k], grads[1][k], rtol=1e-3, atol=2e-5, err_msg=k)
        assert_pgrad_matches_fp64(grads[0][k], g, k)                # fp64 run of the reference class, 1e-4


@pytest.mark.parametrize("train", [True, False])
def test_module_large_batch_forward_path(train):
    """K large enough for the addmm-on-W^T forward (>= 32 M output elements per layer): same outputs as nn.Linear"""
    K = 600_000
    gen = torch.Generator(DEV).manual_seed(11)
    x = torch.randn((K, 4), device=DEV, generator=gen) * torch.tensor([20.0, 20.0, 1.0, 0.3], device=DEV)
    ys = []
    for fused in (True, False):
        torch.manual_seed(3)
        adv = VoxelPerturber(hidden_channels=[64, 128, 64], fused_wgrad=fused, fused_trunk=False, nan_guards=False).build(4, DEV).train(train)
        if train:
            for q in adv.parameters():
                q._backward_hooks = None                       # compare the sums themselves, in front of the +-0.1 clamp
            y, _ = adv(x)
            y.sum().backward()
            ys.append((y.detach(), adv.model[3].weight.grad.clone()))
        else:
            with torch.no_grad():
                ys.append((adv(x)[0], None))
    torch.testing.assert_close(ys[0][0], ys[1][0], rtol=1e-5, atol=1e-5)
    if train:
        # two fp32 summations over 600 k rows (cuBLAS split-K against this library's kernel -- 3 x TF32 on the tensor cores
        # for this 128 x 64 layer, ~5e-7 of scale against fp64): 1e-4 of the tensor's scale
        scale = float(ys[1][1].abs().max())
        torch.testing.assert_close(ys[0][1], ys[1][1], rtol=0, atol=1e-4 * scale)


# ------------------------------------------------------------------ rpc_bn_relu_fwd/bwd (third slice of SURVEY 8f row 2)
@pytest.mark.parametrize("K,H", [(2, 64), (257, 8), (4099, 64), (70001, 128), (1000, 100), (300000, 64)])
def test_bn_relu_vs_fp64(K, H):
    gen = torch.Generator(DEV).manual_seed(K + H)
    y = torch.randn((K, H), device=DEV, generator=gen) * 3 + 1.5
    gamma = torch.randn(H, device=DEV, generator=gen)
    beta = torch.randn(H, device=DEV, generator=gen) * 0.5
    go = torch.randn((K, H), device=DEV, generator=gen)
    rm, rv = torch.zeros(H, device=DEV), torch.ones(H, device=DEV)
    z, mean, invstd = ops.bn_relu_fwd(y, gamma, beta, 1e-3, 0.1, rm, rv)
    y64 = y.double().requires_grad_(True)
    g64, b64 = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    rm64, rv64 = torch.zeros(H, device=DEV, dtype=torch.float64), torch.ones(H, device=DEV, dtype=torch.float64)
    ref = torch.relu(torch.nn.functional.batch_norm(y64, rm64, rv64, g64, b64, True, 0.1, 1e-3))
    torch.testing.assert_close(z.double(), ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(rm.double(), rm64, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(rv.double(), rv64, rtol=1e-5, atol=1e-6)
    # backward where the fp32 and fp64 ReLU masks agree (elements within rounding of the kink may differ)
    mask = ((ref > 1e-4) | (z == 0) & (ref == 0)).all()
    ref.backward(go.double())
    gy, gg, gb = ops.bn_relu_bwd(go, y, mean, invstd, gamma, beta)
    scale = float(y64.grad.abs().max())
    frac_bad = float(((gy.double() - y64.grad).abs() > 1e-4 * scale).double().mean())
    assert frac_bad < 1e-3, frac_bad                             # rows at the kink only
    if bool(mask):
        torch.testing.assert_close(gg.double(), g64.grad, rtol=1e-4, atol=1e-4 * float(g64.grad.abs().max()))
        torch.testing.assert_close(gb.double(), b64.grad, rtol=1e-4, atol=1e-4 * float(b64.grad.abs().max()))


def test_module_fused_bn_equals_aten(golden_dir):
    """VoxelPerturber(fused_bn=True) against nn.BatchNorm1d + nn.ReLU: outputs, parameter gradients, running statistics,
    state-dict keys; eval mode goes through nn.BatchNorm1d itself"""
    g = np.load(os.path.join(golden_dir, "perturber_kitti_train.npz"))
    state = {k[len("state_"):]: torch.from_numpy(v) for k, v in g.items() if k.startswith("state_")}
    x = torch.from_numpy(g["x"]).to(DEV)
    w = torch.from_numpy(g["w"]).to(DEV)
    res = []
    for fused in (True, False):
        adv = VoxelPerturber(hidden_channels=[int(h) for h in g["hidden"]], fused_bn=fused, fused_trunk=False).build(4)
        assert set(adv.state_dict()) == set(state)
        adv.load_state_dict(state)
        adv = adv.to(DEV).train()
        y, ld = adv(x)
        ((y * w).sum() + ld["l2_norm"] + 2 * ld["intensity_loss"] + 3 * ld["bias_loss"] + 4 * ld["imbalance_loss"]).backward()
        adv.eval()
        with torch.no_grad():
            ye = adv(x)[0]
        res.append((y.detach(), {k: q.grad.clone() for k, q in adv.named_parameters()},
                    {k: v.clone() for k, v in adv.state_dict().items() if "running" in k or "tracked" in k}, ye))
    torch.testing.assert_close(res[0][0], res[1][0], rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(res[0][3], res[1][3], rtol=1e-5, atol=1e-6)
    for k in res[0][1]:
        np.testing.assert_allclose(res[0][1][k].cpu().numpy(), res[1][1][k].cpu().numpy(), rtol=1e-3, atol=2e-5, err_msg=k)
        assert_pgrad_matches_fp64(res[0][1][k].cpu().numpy(), g, k)
    for k in res[0][2]:
        torch.testing.assert_close(res[0][2][k].double(), res[1][2][k].double(), rtol=1e-5, atol=1e-6, msg=k)
